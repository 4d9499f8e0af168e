// Fused 2-D Burgers stage kernel for the y-slab program: Lax-Friedrichs split, WENO5-JS face fluxes in x
// and y, flux divergence and the SSP-RK3 stage combine in one pass.
// Reference: Compute_dF + Compute_dG + Reconstruct1d + Compute_RK, MultiGPU/Burgers2d_Baseline/
// Kernels.cu:103-365 (three kernels and an Lu array there).
//
// One CTA owns a TX x TY tile of cells.  (1) u of tile + halo 3 is read once, split into (f+, f-) =
// ({3/4,-1/4} u^2 by sign, exact) and parked as pairs in shared memory.  (2) The (TX+1)*TY x-faces and
// TX*(TY+1) y-faces of the tile are one flat work list spread over all threads -- every face is evaluated
// once per tile (rim faces twice per grid: 2-3 % extra) with weno_side6 of the 3-D ring kernel (one
// reciprocal per side; parity to 1e-12, not bitwise).  (3) Each cell takes the differences of its four
// faces, the RK combine, and is stored (twice with PEER: also into the neighbour GPU's ghost rows).
// Region rules of the reference (Kernels.cu:240,249,273,311,345): rows [own_lo, own_hi) get a right-hand
// side; x: cells i in [2, nx-4] with the flux at 2-1/2 taken as 0; y: columns [3, nx-4]; RK on every cell.
#pragma once
#include "burgers_kernels.cuh"

namespace adv {

template <typename REAL, int TX, int TY>
struct Burgers2dSmem {
    static constexpr int H = 3, RP = TX + 2 * H, ROWS = TY + 2 * H;
    typename V2T<REAL>::type fpm[ROWS][RP];     // (f+, f-) of tile + halo
    REAL fx[TY][TX + 1];                        // 6 x flux at (i0 + c - 1) + 1/2, row r
    REAL fy[TY + 1][TX];                        // 6 x flux at (j0 + r - 1) + 1/2, column c
};

template <typename REAL, int W, int TX, int TY, int NT, int STAGE, bool PEER>
__global__ void __launch_bounds__(NT)
burgers2d_stage_tile(BurgersArgs<REAL> a, const REAL* __restrict__ u, const REAL* __restrict__ uo, REAL* __restrict__ out)
{
    typedef typename V2T<REAL>::type V2;
    typedef Burgers2dSmem<REAL, TX, TY> S;
    constexpr int H = S::H, RP = S::RP, ROWS = S::ROWS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    S& sm = *reinterpret_cast<S*>(smem_raw);
    const int tid = threadIdx.x;
    const int i0 = blockIdx.x * TX, j0 = a.kb + blockIdx.y * TY;      // a.kb/a.ke: row window of this launch
    const long long P = a.pitch;
    const REAL* __restrict__ wc = a.wc;

    // (1) stage (f+, f-) of tile + halo; points outside the array are zero (their faces are never used)
    for (int n = tid; n < ROWS * RP; n += NT) {
        const int r = n / RP, c = n - r * RP;
        const int gi = i0 - H + c, gj = j0 - H + r;
        REAL v = REAL(0);
        if (gi >= 0 && gi < a.nx && gj >= 0 && gj < a.ny) v = u[(long long)gi + P * gj];
        const REAL sq = v * v;
        const bool ng = is_neg<REAL>(v);
        V2 f;
        f.x = ng ? wc[15] * -sq : wc[14] * sq;      // f+ = 0.5*(u^2/2 + |u|u)
        f.y = ng ? wc[14] * sq : wc[15] * -sq;      // f- = 0.5*(u^2/2 - |u|u)
        sm.fpm[r][c] = f;
    }
    __syncthreads();

    // (2) faces.  x-face (r, c): between tile columns c-1 and c of tile row r; y-face (r, c): between tile
    // rows r-1 and r of column c.  In fpm coordinates the cell left of / below the face is at (+H-1).
    constexpr int NFX = TY * (TX + 1), NF = NFX + (TY + 1) * TX;
    for (int n = tid; n < NF; n += NT) {
        if (n < NFX) {
            const int r = n / (TX + 1), c = n - r * (TX + 1);
            const int m = i0 + c - 1, gj = j0 + r;                  // face at m + 1/2
            REAL f6 = REAL(0);
            if (m >= 2 && m <= a.nx - 4 && gj >= a.own_lo && gj < a.own_hi && gj < a.ke) {
                const V2* q = &sm.fpm[r + H][c + H - 1];
                f6 = weno_side6<REAL, W>(q[-2].x, q[-1].x, q[0].x, q[1].x, q[2].x, wc) +
                     weno_side6<REAL, W>(q[3].y, q[2].y, q[1].y, q[0].y, q[-1].y, wc);
            }
            sm.fx[r][c] = f6;
        } else {
            const int t = n - NFX;
            const int r = t / TX, c = t - r * TX;
            const int gi = i0 + c, m = j0 + r - 1;                  // face at m + 1/2
            REAL f6 = REAL(0);
            if (gi >= 3 && gi <= a.nx - 4 && m >= a.own_lo - 1 && m < a.own_hi && m < a.ke) {
                const V2* q = &sm.fpm[r + H - 1][c + H];
                f6 = weno_side6<REAL, W>(q[-2 * RP].x, q[-RP].x, q[0].x, q[RP].x, q[2 * RP].x, wc) +
                     weno_side6<REAL, W>(q[3 * RP].y, q[2 * RP].y, q[RP].y, q[0].y, q[-RP].y, wc);
            }
            sm.fy[r][c] = f6;
        }
    }
    __syncthreads();

    // (3) divergence + RK combine
    for (int n = tid; n < TX * TY; n += NT) {
        const int r = n / TX, c = n - r * TX;
        const int gi = i0 + c, gj = j0 + r;
        if (gi >= a.nx || gj >= a.ke) continue;
        const long long o = (long long)gi + P * gj;
        const REAL v = u[o];
        REAL Lu = REAL(0);
        if (gj >= a.own_lo && gj < a.own_hi && gi >= 2 && gi <= a.nx - 4) {
            Lu = -(sm.fx[r][c + 1] - sm.fx[r][c]) * a.rdx6;         // fx[r][c] is 0 for gi == 2 (m == 1)
            if (gi >= 3) Lu -= (sm.fy[r + 1][c] - sm.fy[r][c]) * a.rdy6;
        }
        REAL qo = v;
        if (STAGE != 1) qo = uo[o];
        const REAL res = rk_stage<REAL, STAGE>(v, qo, Lu, a.dt, wc);
        out[o] = res;
        if (PEER) a.out2[a.off2 + o] = res;
    }
}

}  // namespace adv
