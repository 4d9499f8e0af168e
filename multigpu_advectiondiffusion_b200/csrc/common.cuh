// Shared device/host helpers for libadvdiff (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace adv {

// SSP-RK3 (Shu-Osher) stage combine, fused into every stage kernel.
// Reference: Compute_RK, MultiGPU/Burgers3d_Baseline/Kernels.cu:478-486.
template <typename REAL>
__device__ __forceinline__ REAL rk_combine(int stage, REAL q, REAL qo, REAL Lq, REAL dt)
{
    if (stage == 1) return qo + dt * Lq;
    if (stage == 2) return REAL(0.75) * qo + REAL(0.25) * (q + dt * Lq);
    return (qo + REAL(2) * (q + dt * Lq)) / REAL(3);
}

struct Box3 {           // inclusive index box
    int lo[3], hi[3];
    __host__ __device__ bool has(int i, int j, int k) const
    {
        return i >= lo[0] && i <= hi[0] && j >= lo[1] && j <= hi[1] && k >= lo[2] && k <= hi[2];
    }
};

template <typename REAL>
struct HeatArgs {
    int nx, ny, nzs;
    long long pitch, plane;   // elements
    int kb, ke;               // planes of u_out to produce: [kb, ke)
    Box3 lap;                 // where the Laplacian is evaluated (RHS = 0 elsewhere)
    Box3 rk;                  // where the RK arithmetic runs (u_out = u_in elsewhere)
    int iso;                  // 1: kx/12 * (sum - 90 c)  (MultiGPU/Diffusion3d_Baseline/Kernels.cu:240-243)
    int stage;
    REAL kx, ky, kz, dt;
};

enum { WENO_JS = 0, WENO_Z = 1 };
enum { BMODE_SGPU = 0, BMODE_MGPU = 1 };

template <typename REAL>
struct BurgersArgs {
    int nx, ny, nzs;
    long long pitch, plane;
    int kb, ke;               // planes of u_out to produce
    int own_lo, own_hi;       // planes [own_lo, own_hi) get a RHS
    int stage;
    int visc;                 // SGPU: add kx,ky,kz Laplacian
    int quirk_lo, quirk_hi;   // SGPU: local planes [quirk_lo, quirk_hi] get the viscous term twice (F12); hi<lo: none
    REAL dt, rdx, rdy, rdz;   // spacings passed as given (division kept as in the reference)
    REAL dx, dy, dz;
    REAL kx, ky, kz;
};

}  // namespace adv
