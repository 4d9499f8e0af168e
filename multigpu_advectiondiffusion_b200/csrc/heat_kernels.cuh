// Fused heat-equation stage kernels: 4th-order Laplacian + SSP-RK3 stage combine in one pass.
// Reference operators: Compute_Laplace3d (SingleGPU/Diffusion3d_baselineCode/kernels.cu:66-99),
// Compute_Laplace3d_Async (SingleGPU/Diffusion3d_Blocking/kernels.cu:37-88), LaplaceO4_async
// (MultiGPU/Diffusion3d_Baseline/Kernels.cu:207-261), Compute_Laplace2d
// (SingleGPU/Diffusion2d/kernels.cu:70-88) + Compute_RK of each directory.
#pragma once
#include "common.cuh"

namespace adv {

// ---------------------------------------------------------------------------------------------
// v1 ("direct"): one thread per (i,j) column, marching a chunk of z planes with a 5-deep
// register queue; x/y neighbours come straight from global memory through L1.
// ---------------------------------------------------------------------------------------------
template <typename REAL, int CHUNK>
__global__ void __launch_bounds__(256)
heat3d_stage_direct(HeatArgs<REAL> a, const REAL* __restrict__ u, const REAL* __restrict__ uo,
                    REAL* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k0 = a.kb + blockIdx.z * CHUNK;
    const int k1 = min(k0 + CHUNK, a.ke);
    if (i >= a.nx || j >= a.ny || k0 >= k1) return;

    const long long col = (long long)i + a.pitch * j;
    const bool xy_lap = i >= a.lap.lo[0] && i <= a.lap.hi[0] && j >= a.lap.lo[1] && j <= a.lap.hi[1];
    const bool xy_rk  = i >= a.rk.lo[0] && i <= a.rk.hi[0] && j >= a.rk.lo[1] && j <= a.rk.hi[1];

    auto ld = [&](int k) -> REAL { return (k >= 0 && k < a.nzs) ? u[col + a.plane * k] : REAL(0); };
    REAL below2 = ld(k0 - 2), below = ld(k0 - 1), center = ld(k0), above = ld(k0 + 1), above2;

    for (int k = k0; k < k1; k++) {
        above2 = ld(k + 2);
        const long long o = col + a.plane * k;
        REAL res = center;
        if (xy_rk && k >= a.rk.lo[2] && k <= a.rk.hi[2]) {
            REAL Lu = REAL(0);
            if (xy_lap && k >= a.lap.lo[2] && k <= a.lap.hi[2]) {
                const REAL xm2 = u[o - 2], xm1 = u[o - 1], xp1 = u[o + 1], xp2 = u[o + 2];
                const REAL ym2 = u[o - 2 * a.pitch], ym1 = u[o - a.pitch], yp1 = u[o + a.pitch], yp2 = u[o + 2 * a.pitch];
                if (a.iso) {
                    Lu = a.kx / REAL(12) * (
                        - xm2 + REAL(16) * xm1 + REAL(16) * xp1 - xp2
                        - ym2 + REAL(16) * ym1 - REAL(90) * center + REAL(16) * yp1 - yp2
                        - below2 + REAL(16) * below + REAL(16) * above - above2);
                } else {
                    Lu = a.kx * (-xm2 + REAL(16) * xm1 - REAL(30) * center + REAL(16) * xp1 - xp2) +
                         a.ky * (-ym2 + REAL(16) * ym1 - REAL(30) * center + REAL(16) * yp1 - yp2) +
                         a.kz * (-below2 + REAL(16) * below - REAL(30) * center + REAL(16) * above - above2);
                }
            }
            const REAL qo = (a.stage == 1) ? center : uo[o];
            res = rk_combine<REAL>(a.stage, center, qo, Lu, a.dt);
        }
        out[o] = res;
        below2 = below; below = center; center = above; above = above2;
    }
}

// 2-D: one thread per cell (the 512^2 case is launch/L2 bound, not HBM bound).
template <typename REAL>
__global__ void __launch_bounds__(256)
heat2d_stage_direct(HeatArgs<REAL> a, const REAL* __restrict__ u, const REAL* __restrict__ uo,
                    REAL* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = a.kb + blockIdx.y * blockDim.y + threadIdx.y;       // rows [a.kb, a.ke) of this launch
    if (i >= a.nx || j >= a.ke) return;
    const long long o = (long long)i + a.pitch * j;
    const REAL c = u[o];
    REAL res = c;
    if (i >= a.rk.lo[0] && i <= a.rk.hi[0] && j >= a.rk.lo[1] && j <= a.rk.hi[1]) {
        REAL Lu = REAL(0);
        if (i >= a.lap.lo[0] && i <= a.lap.hi[0] && j >= a.lap.lo[1] && j <= a.lap.hi[1]) {
            Lu = a.kx * (-u[o - 2] + REAL(16) * u[o - 1] - REAL(30) * c + REAL(16) * u[o + 1] - u[o + 2]) +
                 a.ky * (-u[o - 2 * a.pitch] + REAL(16) * u[o - a.pitch] - REAL(30) * c + REAL(16) * u[o + a.pitch] - u[o + 2 * a.pitch]);
        }
        const REAL qo = (a.stage == 1) ? c : uo[o];
        res = rk_combine<REAL>(a.stage, c, qo, Lu, a.dt);
    }
    out[o] = res;
    if (a.out2) a.out2[a.off2 + o] = res;
}

}  // namespace adv

// TMA helpers (cp.async.bulk.tensor + mbarrier) used by the warp-specialised kernel below.
#include <cuda.h>

namespace adv {

// RK combine with the division by 3 of stage 3 written as a multiplication (an IEEE double division
// costs ~25 FP64-pipe instructions per cell and made stage 3 compute-bound: 705 vs 505 us at 512^3)
template <typename REAL, int STAGE>
__device__ __forceinline__ REAL rk_combine_fast(REAL q, REAL qo, REAL Lq, REAL dt)
{
    if (STAGE == 1) return qo + dt * Lq;
    if (STAGE == 2) return REAL(0.75) * qo + REAL(0.25) * (q + dt * Lq);
    return (qo + REAL(2) * (q + dt * Lq)) * REAL(1.0 / 3.0);
}

// 2-D, large grids and y-slabs (MultiGPU/Diffusion2d_Baseline/Kernels.cu:130-188): one thread per PAIR of
// cells (one 16-byte access per row), marching a chunk of rows with a 5-deep register queue in y; the x
// neighbours are the two adjacent aligned pairs, which the neighbouring lanes loaded two rows earlier (L1).
// Rows [a.kb, a.ke) of `out` are produced; with PEER every produced row is also stored to out2 + off2.
template <bool B> struct BoolTag2 { static constexpr bool value = B; };
template <typename REAL> struct Pair2T;
template <> struct Pair2T<double> { typedef double2 type; };
template <> struct Pair2T<float> { typedef float2 type; };

template <typename REAL, int STAGE, bool PEER>
__global__ void __launch_bounds__(128)
heat2d_stage_rows(HeatArgs<REAL> a, int chunk, const REAL* __restrict__ u, const REAL* __restrict__ uo,
                  REAL* __restrict__ out)
{
    typedef typename Pair2T<REAL>::type V2;
    const int i = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    const int j0 = a.kb + blockIdx.y * chunk;
    const int j1 = min(j0 + chunk, a.ke);
    if (i >= a.nx || j0 >= j1) return;
    const bool has1 = i + 1 < a.nx;
    const bool lapx0 = i >= a.lap.lo[0] && i <= a.lap.hi[0], lapx1 = i + 1 >= a.lap.lo[0] && i + 1 <= a.lap.hi[0];
    const bool rkx0 = i >= a.rk.lo[0] && i <= a.rk.hi[0], rkx1 = has1 && i + 1 >= a.rk.lo[0] && i + 1 <= a.rk.hi[0];
    const bool ldl = i >= 2, ldr = (long long)i + 3 < a.pitch;
    const V2 zero = {REAL(0), REAL(0)};
    const REAL* __restrict__ up = u + i;
    auto ld = [&](int j) -> V2 { return (j >= 0 && j < a.ny) ? *reinterpret_cast<const V2*>(up + a.pitch * j) : zero; };
    V2 m2 = ld(j0 - 2), m1 = ld(j0 - 1), c = ld(j0), p1 = ld(j0 + 1), p2;

    // one row; FAST: both cells inside the Laplacian and RK boxes, row j+2 exists -- no tests, so that the
    // unrolled steady-state loop below can issue the loads of several rows back to back
    auto row = [&](int j, auto fast_tag) {
        constexpr bool FAST = decltype(fast_tag)::value;
        const long long o = (long long)i + a.pitch * j;
        p2 = FAST ? *reinterpret_cast<const V2*>(u + o + 2 * a.pitch) : ld(j + 2);
        const bool rky = FAST || (j >= a.rk.lo[1] && j <= a.rk.hi[1]);
        const bool lapy = FAST || (j >= a.lap.lo[1] && j <= a.lap.hi[1]);
        V2 res = c;
        if (FAST || (rky && (rkx0 || rkx1))) {
            V2 Lu = zero;
            if (FAST || (lapy && (lapx0 || lapx1))) {
                const V2 L = (FAST || ldl) ? *reinterpret_cast<const V2*>(u + o - 2) : zero;
                const V2 R = (FAST || ldr) ? *reinterpret_cast<const V2*>(u + o + 2) : zero;
                if (FAST || lapx0) Lu.x = a.kx * (-L.x + REAL(16) * L.y - REAL(30) * c.x + REAL(16) * c.y - R.x) +
                                          a.ky * (-m2.x + REAL(16) * m1.x - REAL(30) * c.x + REAL(16) * p1.x - p2.x);
                if (FAST || lapx1) Lu.y = a.kx * (-L.y + REAL(16) * c.x - REAL(30) * c.y + REAL(16) * R.x - R.y) +
                                          a.ky * (-m2.y + REAL(16) * m1.y - REAL(30) * c.y + REAL(16) * p1.y - p2.y);
            }
            V2 qo = c;
            if (STAGE != 1) qo = *reinterpret_cast<const V2*>(uo + o);
            if (FAST || rkx0) res.x = rk_combine_fast<REAL, STAGE>(c.x, qo.x, Lu.x, a.dt);
            if (FAST || rkx1) res.y = rk_combine_fast<REAL, STAGE>(c.y, qo.y, Lu.y, a.dt);
        }
        if (FAST || has1) {
            *reinterpret_cast<V2*>(out + o) = res;
            if (PEER) *reinterpret_cast<V2*>(a.out2 + a.off2 + o) = res;
        } else {
            out[o] = res.x;
            if (PEER) a.out2[a.off2 + o] = res.x;
        }
        m2 = m1; m1 = c; c = p1; p1 = p2;
    };
    const bool fast_x = lapx0 && lapx1 && rkx0 && rkx1 && ldl && ldr;
    const int f0 = fast_x ? max(j0, max(a.lap.lo[1], a.rk.lo[1])) : j1;
    const int f1 = fast_x ? min(j1, min(min(a.lap.hi[1], a.rk.hi[1]) + 1, a.ny - 2)) : j1;      // fast rows [f0, f1)
    int j = j0;
    for (; j < min(f0, j1); j++) row(j, BoolTag2<false>());
#pragma unroll 4
    for (; j < f1; j++) row(j, BoolTag2<true>());
    for (; j < j1; j++) row(j, BoolTag2<false>());
}

template <typename REAL> struct Vec2T;
template <> struct Vec2T<double> { typedef double2 type; };
template <> struct Vec2T<float> { typedef float2 type; };

__device__ __forceinline__ void tma_load_3d(unsigned smem_dst, const CUtensorMap* tmap, int x, int y, int z, unsigned bar)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(smem_dst), "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes)
{
    asm volatile("{\n.reg .b64 st;\nmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_init_(unsigned bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait_parity(unsigned bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "HEAT_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra HEAT_DONE;\n"
        "bra HEAT_WAIT;\n"
        "HEAT_DONE:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}

// ---------------------------------------------------------------------------------------------
// heat3d_stage_tma: TMA-staged tiles, 2.5-D streaming along z, as a producer/consumer pipeline.
//   A CTA owns a 64 x TY tile (two x-adjacent cells per consumer thread) and marches a chunk of z planes.
//   - one producer warp (a single elected lane) keeps NST planes in flight: per plane ONE 3-D TMA box
//     of u (tile + halo 2) and, for stages 2 and 3, one box of uo (tile only), both completing on the
//     stage's `full` mbarrier; it refills a slot as soon as the consumers' `empty` mbarrier fires;
//   - TY consumer warps never meet at a CTA barrier: wait(full) -> stencil -> arrive(empty);
//   - the 5-plane z queue rotates by register renaming (loop unrolled 5x), no moves;
//   - tiles inside the update region run without per-cell masks.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_arrive_(unsigned bar)
{
    asm volatile("{\n.reg .b64 st;\nmbarrier.arrive.shared::cta.b64 st, [%0];\n}" ::"r"(bar) : "memory");
}

// TWO_D: the same pipeline for a 2-D field seen as nx x 1 x rows -- the marching axis is the 2-D y, the tile has
// one row (TY = 1) and no y halo, the Laplacian is kx*X + kz*Z with the 2-D ky passed as kz (same operand order
// as heat2d_stage_rows, so the two kernels give the same bits).
template <typename REAL, int TY, int NST, bool TWO_D = false>
struct HeatTmaSmem {
    static constexpr int TX = 64, H = 2;
    static constexpr int HY = TWO_D ? 0 : H;
    static constexpr int HX = (sizeof(REAL) == 8) ? 2 : 4;
    static constexpr int BX = TX + 2 * HX, BY = TY + 2 * HY;
    struct alignas(128) Plane { REAL v[BY][BX]; };
    struct alignas(128) Tile { REAL v[TY][TX]; };
    Plane stage[NST];
    Tile old[NST];
    unsigned long long full[NST], empty[NST];
};

template <typename REAL, int TY, int NST, int STAGE, bool TWO_D = false>
__global__ void __launch_bounds__(32 * (TY + 1))
heat3d_stage_tma(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_uo,
                  HeatArgs<REAL> a, SlabSched sched, REAL* __restrict__ out)
{
    typedef typename Vec2T<REAL>::type V2;
    typedef HeatTmaSmem<REAL, TY, NST, TWO_D> S;
    constexpr int TX = S::TX, H = S::HY, HX = S::HX, BX = S::BX, BY = S::BY;
    constexpr unsigned kBytesU = BX * BY * sizeof(REAL), kBytesO = TX * TY * sizeof(REAL);
    extern __shared__ __align__(128) unsigned char heat_smem_raw[];
    S& sm = *reinterpret_cast<S*>(heat_smem_raw);

    const int tx = threadIdx.x, ty = threadIdx.y;
    const int i0 = blockIdx.x * TX, j0 = blockIdx.y * TY;
    int kb, ke;
    const int bw = slab_window(sched, kb, ke);          // boundary window (fused halo exchange) or interior chunk
    if (kb >= ke) return;
    REAL* const out2 = (REAL*)slab_out2(sched, bw, a.out2);
    const long long off2 = slab_off2(sched, bw, a.off2);
    const int p_first = kb - 2;
    const int N = (ke + 1) - p_first + 1;                   // planes to stream: kb-2 .. ke+1

    if (tx == 0 && ty == 0) {
        for (int s = 0; s < NST; s++) {
            mbar_init_((unsigned)__cvta_generic_to_shared(&sm.full[s]), 1);
            mbar_init_((unsigned)__cvta_generic_to_shared(&sm.empty[s]), TY);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    slab_wait(sched, bw, tx == 0 && ty == 0);           // the neighbour's previous stage has landed in our ghost planes

    if (ty == TY) {
        // ---------------- producer warp ----------------
        if (tx == 0) {
            for (int n = 0; n < N; n++) {
                const int s = n % NST;
                if (n >= NST) mbar_wait_parity((unsigned)__cvta_generic_to_shared(&sm.empty[s]), ((n / NST) - 1) & 1);
                const unsigned bar = (unsigned)__cvta_generic_to_shared(&sm.full[s]);
                const int p = p_first + n, k = p - 2;
                const bool with_old = (STAGE > 1) && (k >= kb) && (k < ke);
                mbar_expect_tx(bar, with_old ? kBytesU + kBytesO : kBytesU);
                tma_load_3d((unsigned)__cvta_generic_to_shared(&sm.stage[s].v[0][0]), &tmap, i0 - HX, j0 - H, p, bar);
                if (with_old) tma_load_3d((unsigned)__cvta_generic_to_shared(&sm.old[s].v[0][0]), &tmap_uo, i0, j0, k, bar);
            }
        }
        slab_signal(sched, false);                  // meets the consumer warps at the CTA barrier inside
        return;
    }

    // ---------------- consumer warps ----------------
    const int i = i0 + 2 * tx, j = j0 + ty;
    const bool in_dom = (i < a.nx) && (j < a.ny);
    const long long col = (long long)i + a.pitch * j;
    const bool interior = i0 >= a.lap.lo[0] && i0 + TX - 1 <= a.lap.hi[0] && j0 >= a.lap.lo[1] && j0 + TY - 1 <= a.lap.hi[1] &&
                          i0 >= a.rk.lo[0] && i0 + TX - 1 <= a.rk.hi[0] && j0 >= a.rk.lo[1] && j0 + TY - 1 <= a.rk.hi[1];
    const bool xl0 = i >= a.lap.lo[0] && i <= a.lap.hi[0], xl1 = i + 1 >= a.lap.lo[0] && i + 1 <= a.lap.hi[0];
    const bool yl = j >= a.lap.lo[1] && j <= a.lap.hi[1];
    const bool xr0 = i >= a.rk.lo[0] && i <= a.rk.hi[0], xr1 = i + 1 >= a.rk.lo[0] && i + 1 <= a.rk.hi[0];
    const bool yr = j >= a.rk.lo[1] && j <= a.rk.hi[1];
    const REAL ciso = a.kx / REAL(12);
    REAL* po = out + col + a.plane * (kb - 1);              // bumped before each store: plane k

    auto lap = [&](const V2& zm2, const V2& zm1, const V2& zc, const V2& zp1, const V2& zp2, const REAL* row, V2& Lu) {
        const V2 xl = *reinterpret_cast<const V2*>(row - 2), xr = *reinterpret_cast<const V2*>(row + 2);
        if (TWO_D) {
            Lu.x = a.kx * (-xl.x + REAL(16) * xl.y - REAL(30) * zc.x + REAL(16) * zc.y - xr.x) +
                   a.kz * (-zm2.x + REAL(16) * zm1.x - REAL(30) * zc.x + REAL(16) * zp1.x - zp2.x);
            Lu.y = a.kx * (-xl.y + REAL(16) * zc.x - REAL(30) * zc.y + REAL(16) * xr.x - xr.y) +
                   a.kz * (-zm2.y + REAL(16) * zm1.y - REAL(30) * zc.y + REAL(16) * zp1.y - zp2.y);
            return;
        }
        const V2 ym2 = *reinterpret_cast<const V2*>(row - 2 * BX), ym1 = *reinterpret_cast<const V2*>(row - BX);
        const V2 yp1 = *reinterpret_cast<const V2*>(row + BX), yp2 = *reinterpret_cast<const V2*>(row + 2 * BX);
        if (a.iso) {
            Lu.x = ciso * (-xl.x + REAL(16) * xl.y + REAL(16) * zc.y - xr.x
                           - ym2.x + REAL(16) * ym1.x - REAL(90) * zc.x + REAL(16) * yp1.x - yp2.x
                           - zm2.x + REAL(16) * zm1.x + REAL(16) * zp1.x - zp2.x);
            Lu.y = ciso * (-xl.y + REAL(16) * zc.x + REAL(16) * xr.x - xr.y
                           - ym2.y + REAL(16) * ym1.y - REAL(90) * zc.y + REAL(16) * yp1.y - yp2.y
                           - zm2.y + REAL(16) * zm1.y + REAL(16) * zp1.y - zp2.y);
        } else {
            Lu.x = a.kx * (-xl.x + REAL(16) * xl.y - REAL(30) * zc.x + REAL(16) * zc.y - xr.x) +
                   a.ky * (-ym2.x + REAL(16) * ym1.x - REAL(30) * zc.x + REAL(16) * yp1.x - yp2.x) +
                   a.kz * (-zm2.x + REAL(16) * zm1.x - REAL(30) * zc.x + REAL(16) * zp1.x - zp2.x);
            Lu.y = a.kx * (-xl.y + REAL(16) * zc.x - REAL(30) * zc.y + REAL(16) * xr.x - xr.y) +
                   a.ky * (-ym2.y + REAL(16) * ym1.y - REAL(30) * zc.y + REAL(16) * yp1.y - yp2.y) +
                   a.kz * (-zm2.y + REAL(16) * zm1.y - REAL(30) * zc.y + REAL(16) * zp1.y - zp2.y);
        }
    };

    int n = 0;
    // one plane: `nw` receives plane p (it held the dead plane p-5); stencil planes are zm2..zp1, nw
    auto step = [&](V2& nw, const V2& zm2, const V2& zm1, const V2& zc, const V2& zp1) -> bool {
        if (n >= N) return false;
        const int s = n % NST;
        mbar_wait_parity((unsigned)__cvta_generic_to_shared(&sm.full[s]), (n / NST) & 1);
        nw = *reinterpret_cast<const V2*>(&sm.stage[s].v[ty + H][2 * tx + HX]);
        const int k = p_first + n - 2;
        if (k >= kb) {
            const int sk = (n - 2) % NST;
            po += a.plane;
            const bool zl = k >= a.lap.lo[2] && k <= a.lap.hi[2];
            const bool zr = k >= a.rk.lo[2] && k <= a.rk.hi[2];
            V2 qo = zc;
            if (STAGE > 1) qo = *reinterpret_cast<const V2*>(&sm.old[s].v[ty][2 * tx]);
            V2 res = zc;
            const REAL* row = &sm.stage[sk].v[ty + H][2 * tx + HX];
            if (interior) {
                if (zr) {
                    V2 Lu;
                    Lu.x = Lu.y = REAL(0);
                    if (zl) lap(zm2, zm1, zc, zp1, nw, row, Lu);
                    res.x = rk_combine_fast<REAL, STAGE>(zc.x, qo.x, Lu.x, a.dt);
                    res.y = rk_combine_fast<REAL, STAGE>(zc.y, qo.y, Lu.y, a.dt);
                }
                *reinterpret_cast<V2*>(po) = res;
                if (out2) *reinterpret_cast<V2*>(out2 + (po - out) + off2) = res;
            } else {
                if (yr && zr) {
                    V2 Lu;
                    Lu.x = Lu.y = REAL(0);
                    if (yl && zl) {
                        lap(zm2, zm1, zc, zp1, nw, row, Lu);
                        Lu.x = xl0 ? Lu.x : REAL(0);
                        Lu.y = xl1 ? Lu.y : REAL(0);
                    }
                    const REAL r0 = rk_combine_fast<REAL, STAGE>(zc.x, qo.x, Lu.x, a.dt);
                    const REAL r1 = rk_combine_fast<REAL, STAGE>(zc.y, qo.y, Lu.y, a.dt);
                    res.x = xr0 ? r0 : zc.x;
                    res.y = xr1 ? r1 : zc.y;
                }
                if (in_dom) {
                    *reinterpret_cast<V2*>(po) = res;
                    if (out2) *reinterpret_cast<V2*>(out2 + (po - out) + off2) = res;
                }
            }
        }
        if (n >= 2) {
            // the plane staged two iterations ago has served as centre plane (or never will: the two
            // run-in planes below the chunk): its slot goes back to the producer
            __syncwarp();
            if (tx == 0) mbar_arrive_((unsigned)__cvta_generic_to_shared(&sm.empty[(n - 2) % NST]));
        }
        n++;
        return true;
    };

    V2 q0, q1, q2, q3, q4;
    q0.x = q0.y = q1.x = q1.y = q2.x = q2.y = q3.x = q3.y = q4.x = q4.y = REAL(0);
    for (;;) {
        if (!step(q0, q1, q2, q3, q4)) break;
        if (!step(q1, q2, q3, q4, q0)) break;
        if (!step(q2, q3, q4, q0, q1)) break;
        if (!step(q3, q4, q0, q1, q2)) break;
        if (!step(q4, q0, q1, q2, q3)) break;
    }
    slab_signal(sched, tx == 0 && ty == 0);         // last CTA of a boundary window raises the neighbour's flag
}

}  // namespace adv
