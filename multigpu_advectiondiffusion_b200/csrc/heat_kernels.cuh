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
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= a.nx || j >= a.ny) return;
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
}

}  // namespace adv
