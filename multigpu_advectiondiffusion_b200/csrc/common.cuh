// Shared device/host helpers for libadvdiff (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>

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
    // fused halo exchange: when non-null every produced plane is ALSO stored to out2 + off2 (a neighbour
    // GPU's ghost planes, mapped through NVLink peer access); off2 is an element offset
    REAL* out2;
    long long off2;
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
    REAL rdx6, rdy6, rdz6;    // 1/(6 d): the ring5 kernel carries six times the face flux (saves the /6 per side)
    REAL dx, dy, dz;
    REAL kx, ky, kz;
    REAL* out2;               // fused halo exchange: second destination of every produced plane (peer ghost planes)
    long long off2;
    // WENO constants read straight from the kernel-parameter constant bank (no immediate moves):
    // [0..2] r_k*13/12, [3..5] r_k/4, [6..8] r_k*eps with r_k = 1/sqrt(d_k) (JS) or 1 (Z);
    // [9] 2/6 [10] 7/6 [11] 11/6 [12] 1/6 [13] 5/6 [14] 0.75 [15] 0.25 [16..18] d_k = .1 .6 .3 [19] 1/3
    REAL wc[20];
    // the pair kernel's copy of [0..8] with the k = 0 entries divided by sqrt(2) (JS): P_0 arrives halved, which
    // turns 2*d01*n0 + d12*n2 over n0+n1+n2 into d01*n0 + d12*n2' over n0/2 + n1' + n2' (one addition less)
    REAL wp[9];
    // dynamic CFL step (adv_run_until_cfl): when cfl_done is non-null the kernel returns at once if *cfl_done,
    // takes dt from *cfl_dt instead of the `dt` argument, and (stage 3, cfl_umax non-null) folds max|u_out| of
    // everything it stores into *cfl_umax (IEEE bits of a non-negative double order like the values)
    const double* cfl_dt;
    const int* cfl_done;
    unsigned long long* cfl_umax;
};

template <typename REAL>
inline void fill_weno_consts(BurgersArgs<REAL>& b, int weno)
{
    const double d[3] = {0.1, 0.6, 0.3};
    for (int k = 0; k < 3; k++) {
        const double r = weno == WENO_JS ? 1.0 / sqrt(d[k]) : 1.0;
        b.wc[k] = (REAL)(r * 13.0 / 12.0);
        b.wc[3 + k] = (REAL)(r * 0.25);
        b.wc[6 + k] = (REAL)(r * 1e-6);
        b.wc[16 + k] = (REAL)d[k];
    }
    b.wc[9] = (REAL)(2.0 / 6); b.wc[10] = (REAL)(7.0 / 6); b.wc[11] = (REAL)(11.0 / 6);
    b.wc[12] = (REAL)(1.0 / 6); b.wc[13] = (REAL)(5.0 / 6); b.wc[14] = (REAL)0.75; b.wc[15] = (REAL)0.25;
    b.wc[19] = (REAL)(1.0 / 3.0);
    for (int k = 0; k < 3; k++) {
        const double r = weno == WENO_JS ? 1.0 / sqrt(d[k]) : 1.0;
        const double h = (weno == WENO_JS && k == 0) ? 1.0 / sqrt(2.0) : 1.0;
        b.wp[k] = (REAL)(h * r * 13.0 / 12.0);
        b.wp[3 + k] = (REAL)(h * r * 0.25);
        b.wp[6 + k] = (REAL)(h * r * 1e-6);
    }
}

}  // namespace adv
