// Fused 2-D Burgers stage kernel for the y-slab program: Lax-Friedrichs split, WENO5-JS face fluxes in x
// and y, flux divergence and the SSP-RK3 stage combine in one pass.
// Reference: Compute_dF + Compute_dG + Reconstruct1d + Compute_RK, MultiGPU/Burgers2d_Baseline/
// Kernels.cu:103-365 (three kernels and an Lu array there).
//
// One CTA of 256 threads owns a 64 x 16 tile of cells; thread (cx, ry) owns cells (ry + 4p, cx), p = 0..3.
// (1) u of tile + halo 3 is read once, split into (f+, f-) = ({3/4,-1/4} u^2 by sign, exact) and parked as
//     pairs in shared memory; a thread keeps its own four u values in registers.
// (2) Faces: thread (cx, ry) evaluates the x-face left of and the y-face below each of its cells (eight
//     compile-time-unrolled evaluations, no index arithmetic), and one extra pass of three warps does the
//     rim (the 16 x-faces right of the tile, the 64 y-faces above it).  Every face is evaluated once per
//     tile with weno_side6 of the 3-D ring kernel (one reciprocal per side; parity to 1e-12, not bitwise).
// (3) Each cell takes the differences of its four faces (two of them its own registers), the RK combine,
//     and is stored (twice with PEER: also into the neighbour GPU's ghost rows).
// Tiles whose faces and cells all lie inside the update region run a body compiled without any of the
// region tests (FAST).  Region rules of the reference (Kernels.cu:240,249,273,311,345): rows
// [own_lo, own_hi) get a right-hand side; x: cells i in [2, nx-4] with the flux at 2-1/2 taken as 0;
// y: columns [3, nx-4]; RK on every cell.
#pragma once
#include "burgers_kernels.cuh"

namespace adv {

// Parity anchor (ADV_BURGERS_DIRECT=1), the 2-D twin of burgers3d_stage_direct: one thread per cell, every
// thread evaluates its own four faces with the reference's operand order and IEEE divisions
// (Reconstruct1d, Kernels.cu:103-212; `Lu = -(fu-fu_old)/dx`, `Lu -= (fu-fu_old)/dy`, :273,:345).  Slow on
// purpose; it separates "different rounding" from "different result" when the tile kernel is questioned.
template <typename REAL, int W, bool PEER>
__global__ void __launch_bounds__(256)
burgers2d_stage_direct(BurgersArgs<REAL> a, const REAL* __restrict__ u, const REAL* __restrict__ uo, REAL* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = a.kb + blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= a.nx || j >= a.ke) return;
    const long long P = a.pitch, o = (long long)i + P * j;
    const REAL q = u[o];
    REAL Lu = REAL(0);
    if (j >= a.own_lo && j < a.own_hi) {
        if (i >= 2 && i <= a.nx - 4) {          // x sweep: cells 2..nx-4, flux at 2-1/2 taken as 0 (Kernels.cu:249,273)
            const REAL fp = face_flux_ref<REAL, W>(u[o - 2], u[o - 1], q, u[o + 1], u[o + 2], u[o + 3]);
            const REAL fm = (i == 2) ? REAL(0) : face_flux_ref<REAL, W>(u[o - 3], u[o - 2], u[o - 1], q, u[o + 1], u[o + 2]);
            Lu = -(fp - fm) / a.dx;
        }
        if (i >= 3 && i <= a.nx - 4) {          // y sweep: columns 3..nx-4, real seed flux (Kernels.cu:311-345)
            const REAL gp = face_flux_ref<REAL, W>(u[o - 2 * P], u[o - P], q, u[o + P], u[o + 2 * P], u[o + 3 * P]);
            const REAL gm = face_flux_ref<REAL, W>(u[o - 3 * P], u[o - 2 * P], u[o - P], q, u[o + P], u[o + 2 * P]);
            Lu -= (gp - gm) / a.dy;
        }
    }
    const REAL qo = (a.stage == 1) ? q : uo[o];
    const REAL res = rk_combine<REAL>(a.stage, q, qo, Lu, a.dt);
    out[o] = res;
    if (PEER) a.out2[a.off2 + o] = res;
}

template <typename REAL>
struct Burgers2dSmem {
    static constexpr int TX = 64, TY = 16, H = 3, RP = TX + 2 * H, ROWS = TY + 2 * H;
    typename V2T<REAL>::type fpm[ROWS][RP];     // (f+, f-) of tile + halo
    REAL fx[TY][TX + 1];                        // 6 x flux at (i0 + c - 1) + 1/2, row r
    REAL fy[TY + 1][TX];                        // 6 x flux at (j0 + r - 1) + 1/2, column c
};

template <typename REAL, int W, int STAGE, bool PEER>
__global__ void __launch_bounds__(256, 4)
burgers2d_stage_tile(BurgersArgs<REAL> a, const REAL* __restrict__ u, const REAL* __restrict__ uo, REAL* __restrict__ out)
{
    typedef typename V2T<REAL>::type V2;
    typedef Burgers2dSmem<REAL> S;
    constexpr int TX = S::TX, TY = S::TY, H = S::H, RP = S::RP, NT = 256, NP = TY / 4;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    S& sm = *reinterpret_cast<S*>(smem_raw);
    const int tid = threadIdx.x, cx = tid & (TX - 1), ry = tid >> 6;
    const int i0 = blockIdx.x * TX, j0 = a.kb + blockIdx.y * TY;      // a.kb/a.ke: row window of this launch
    const long long P = a.pitch;
    const REAL* __restrict__ wc = a.wc;
    const int gi = i0 + cx;

    auto split = [&](REAL v) -> V2 {
        const REAL sq = v * v;
        const bool ng = is_neg<REAL>(v);
        V2 f;
        f.x = ng ? wc[15] * -sq : wc[14] * sq;      // f+ = 0.5*(u^2/2 + |u|u)
        f.y = ng ? wc[14] * sq : wc[15] * -sq;      // f- = 0.5*(u^2/2 - |u|u)
        return f;
    };
    // face at (cell m) + 1/2 along a line of the (f+, f-) tile with element stride s; q points at cell m
    // round 2: the 40-instruction side of the 3-D pair kernel (weno_core6: forward differences, halved P_0, one cubic
    // reciprocal step) instead of weno_side6 (43)
    const REAL* __restrict__ we = (W == WENO_JS) ? a.wp : a.wc;
    auto face6 = [&](const V2* q, int s) -> REAL {
        const V2 qa = q[-2 * s], qb = q[-s], qc = q[0], qd = q[s], qe = q[2 * s], qf = q[3 * s];
        const REAL g0 = qb.x - qa.x, g1 = qc.x - qb.x, g2 = qd.x - qc.x, g3 = qe.x - qd.x;
        const REAL hn = weno_core6<REAL, W, false>(qc.x, g0, g1, g2, g3, g1 - g0, g2 - g1, g3 - g2, we, wc);      // centred on c, from f+
        const REAL h0 = qc.y - qb.y, h1 = qd.y - qc.y, h2 = qe.y - qd.y, h3 = qf.y - qe.y;
        const REAL hp = weno_core6<REAL, W, true>(qd.y, h0, h1, h2, h3, h1 - h0, h2 - h1, h3 - h2, we, wc);       // centred on d, from f-, read backwards
        return hn + hp;
    };

    // (1) own cells (kept in registers), then the halo frame; points outside the array are zero
    REAL v[NP];
    const bool col_in = gi < a.nx;
#pragma unroll
    for (int p = 0; p < NP; p++) {
        const int r = ry + 4 * p, gj = j0 + r;
        v[p] = (col_in && gj < a.ny) ? u[(long long)gi + P * gj] : REAL(0);
        sm.fpm[r + H][cx + H] = split(v[p]);
    }
    constexpr int NTB = 2 * H * RP, NHALO = NTB + 2 * H * TY;      // 3 rows below, 3 above, 3+3 columns beside
    for (int n = tid; n < NHALO; n += NT) {
        int r, c;
        if (n < NTB) { r = n / RP; c = n - r * RP; if (r >= H) r += TY; }
        else { const int t = n - NTB; r = H + t / (2 * H); c = t - (t / (2 * H)) * (2 * H); if (c >= H) c += TX; }
        const int hi = i0 - H + c, hj = j0 - H + r;
        REAL hv = REAL(0);
        if (hi >= 0 && hi < a.nx && hj >= 0 && hj < a.ny) hv = u[(long long)hi + P * hj];
        sm.fpm[r][c] = split(hv);
    }
    __syncthreads();

    // tiles whose 65 x 16 x-faces, 64 x 17 y-faces and 64 x 16 cells are all inside the update region
    const int jhi = min(a.own_hi, a.ke);
    const bool fast = i0 - 1 >= 3 && i0 + TX <= a.nx - 4 && j0 - 1 >= a.own_lo && j0 + TY <= jhi;

    auto body = [&](auto fast_tag) {
        constexpr bool FAST = decltype(fast_tag)::value;
        // (2) own faces: left of and below each own cell
        REAL fxl[NP], fyb[NP];
#pragma unroll
        for (int p = 0; p < NP; p++) {
            const int r = ry + 4 * p, gj = j0 + r;
            const int mx = gi - 1, my = gj - 1;
            fxl[p] = REAL(0); fyb[p] = REAL(0);
            if (FAST || (mx >= 2 && mx <= a.nx - 4 && gj >= a.own_lo && gj < jhi))
                fxl[p] = face6(&sm.fpm[r + H][cx + H - 1], 1);
            if (FAST || (gi >= 3 && gi <= a.nx - 4 && my >= a.own_lo - 1 && my < jhi))
                fyb[p] = face6(&sm.fpm[r + H - 1][cx + H], RP);
            sm.fx[r][cx] = fxl[p];
            sm.fy[r][cx] = fyb[p];
        }
        // rim: warp 0 lanes 0..15 the x-faces right of the tile, warps 1..2 the y-faces above it
        if (tid < TY) {
            const int r = tid, gj = j0 + r, mx = i0 + TX - 1;
            REAL f = REAL(0);
            if (FAST || (mx >= 2 && mx <= a.nx - 4 && gj >= a.own_lo && gj < jhi)) f = face6(&sm.fpm[r + H][TX + H - 1], 1);
            sm.fx[r][TX] = f;
        } else if (tid >= 32 && tid < 32 + TX) {
            const int c = tid - 32, hi = i0 + c, my = j0 + TY - 1;
            REAL f = REAL(0);
            if (FAST || (hi >= 3 && hi <= a.nx - 4 && my >= a.own_lo - 1 && my < jhi)) f = face6(&sm.fpm[TY + H - 1][c + H], RP);
            sm.fy[TY][c] = f;
        }
        __syncthreads();
        // (3) divergence + RK combine
#pragma unroll
        for (int p = 0; p < NP; p++) {
            const int r = ry + 4 * p, gj = j0 + r;
            if (!FAST && (!col_in || gj >= a.ke)) continue;
            const long long o = (long long)gi + P * gj;
            REAL Lu = REAL(0);
            if (FAST || (gj >= a.own_lo && gj < a.own_hi && gi >= 2 && gi <= a.nx - 4)) {
                Lu = -(sm.fx[r][cx + 1] - fxl[p]) * a.rdx6;            // fxl is 0 for gi == 2 (face 1 + 1/2)
                if (FAST || gi >= 3) Lu -= (sm.fy[r + 1][cx] - fyb[p]) * a.rdy6;
            }
            REAL qo = v[p];
            if (STAGE != 1) qo = uo[o];
            const REAL res = rk_stage<REAL, STAGE>(v[p], qo, Lu, a.dt, wc);
            out[o] = res;
            if (PEER) a.out2[a.off2 + o] = res;
        }
    };
    if (fast) body(BoolTag<true>()); else body(BoolTag<false>());
}

}  // namespace adv
