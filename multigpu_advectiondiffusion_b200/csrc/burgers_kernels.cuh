// Fused 3-D Burgers stage kernels: Lax-Friedrichs split, WENO5 (JS or Z) face fluxes in x, y, z,
// flux divergence, optional viscous 13-point term and the SSP-RK3 stage combine in ONE pass.
// Reference: Compute_dF/dG/dH + Reconstruct1d + Compute_RK, MultiGPU/Burgers3d_Baseline/
// Kernels.cu:112-486; WENO5Zreconstruction + Compute_Laplace_Async + Compute_RK,
// SingleGPU/Burgers3d_WENO5/kernels.cu:142-530.
#pragma once
#include "common.cuh"

namespace adv {

// ---- Lax-Friedrichs split with pointwise a=|u| (Kernels.cu:32-35,256-264) -------------------
template <typename REAL> __device__ __forceinline__ REAL lf_plus(REAL u)
{
    return REAL(0.5) * (REAL(0.5) * u * u + fabs(u) * u);
}
template <typename REAL> __device__ __forceinline__ REAL lf_minus(REAL u)
{
    return REAL(0.5) * (REAL(0.5) * u * u - fabs(u) * u);
}

// ---- WENO5 one-sided reconstructions, operand order of the reference ------------------------
// minus side: f^- at the face from f+ on cells (c-2..c+2); linear weights (.1,.6,.3)
template <typename REAL, int W>
__device__ __forceinline__ REAL weno_minus_ref(REAL vmm, REAL vm, REAL v, REAL vp, REAL vpp)
{
    const REAL C1312 = REAL(13.0 / 12.0), C14 = REAL(0.25), EPS = REAL(1e-6);
    const REAL B0 = C1312 * (vmm - 2 * vm + v) * (vmm - 2 * vm + v) + C14 * (vmm - 4 * vm + 3 * v) * (vmm - 4 * vm + 3 * v);
    const REAL B1 = C1312 * (vm - 2 * v + vp) * (vm - 2 * v + vp) + C14 * (vm - vp) * (vm - vp);
    const REAL B2 = C1312 * (v - 2 * vp + vpp) * (v - 2 * vp + vpp) + C14 * (3 * v - 4 * vp + vpp) * (3 * v - 4 * vp + vpp);
    const REAL q0 = 2 * vmm - 7 * vm + 11 * v, q1 = -vm + 5 * v + 2 * vp, q2 = 2 * v + 5 * vp - vpp;
    if (W == WENO_JS) {
        const REAL a0 = REAL(0.1) / ((EPS + B0) * (EPS + B0));
        const REAL a1 = REAL(0.6) / ((EPS + B1) * (EPS + B1));
        const REAL a2 = REAL(0.3) / ((EPS + B2) * (EPS + B2));
        const REAL s = a0 + a1 + a2;
        return ((a0 / s) * q0 + (a1 / s) * q1 + (a2 / s) * q2) / 6;
    } else {
        const REAL tau5 = fabs(B0 - B2);
        const REAL a0 = REAL(0.1) * (REAL(1) + tau5 / (B0 + EPS));
        const REAL a1 = REAL(0.6) * (REAL(1) + tau5 / (B1 + EPS));
        const REAL a2 = REAL(0.3) * (REAL(1) + tau5 / (B2 + EPS));
        const REAL s = a0 + a1 + a2;
        return (a0 * q0 + a1 * q1 + a2 * q2) / (6 * s);
    }
}
// plus side: f^+ at the face from f- on cells (c-1..c+3); linear weights (.3,.6,.1)
template <typename REAL, int W>
__device__ __forceinline__ REAL weno_plus_ref(REAL umm, REAL um, REAL u, REAL up, REAL upp)
{
    const REAL C1312 = REAL(13.0 / 12.0), C14 = REAL(0.25), EPS = REAL(1e-6);
    const REAL B0 = C1312 * (umm - 2 * um + u) * (umm - 2 * um + u) + C14 * (umm - 4 * um + 3 * u) * (umm - 4 * um + 3 * u);
    const REAL B1 = C1312 * (um - 2 * u + up) * (um - 2 * u + up) + C14 * (um - up) * (um - up);
    const REAL B2 = C1312 * (u - 2 * up + upp) * (u - 2 * up + upp) + C14 * (3 * u - 4 * up + upp) * (3 * u - 4 * up + upp);
    const REAL q0 = -umm + 5 * um + 2 * u, q1 = 2 * um + 5 * u - up, q2 = 11 * u - 7 * up + 2 * upp;
    if (W == WENO_JS) {
        const REAL a0 = REAL(0.3) / ((EPS + B0) * (EPS + B0));
        const REAL a1 = REAL(0.6) / ((EPS + B1) * (EPS + B1));
        const REAL a2 = REAL(0.1) / ((EPS + B2) * (EPS + B2));
        const REAL s = a0 + a1 + a2;
        return ((a0 / s) * q0 + (a1 / s) * q1 + (a2 / s) * q2) / 6;
    } else {
        const REAL tau5 = fabs(B0 - B2);
        const REAL a0 = REAL(0.3) * (REAL(1) + tau5 / (B0 + EPS));
        const REAL a1 = REAL(0.6) * (REAL(1) + tau5 / (B1 + EPS));
        const REAL a2 = REAL(0.1) * (REAL(1) + tau5 / (B2 + EPS));
        const REAL s = a0 + a1 + a2;
        return (a0 * q0 + a1 * q1 + a2 * q2) / (6 * s);
    }
}

// face flux at c+1/2 from the six values u[c-2..c+3]
template <typename REAL, int W>
__device__ __forceinline__ REAL face_flux_ref(REAL um2, REAL um1, REAL u0, REAL up1, REAL up2, REAL up3)
{
    const REAL hn = weno_minus_ref<REAL, W>(lf_plus(um2), lf_plus(um1), lf_plus(u0), lf_plus(up1), lf_plus(up2));
    const REAL hp = weno_plus_ref<REAL, W>(lf_minus(um1), lf_minus(u0), lf_minus(up1), lf_minus(up2), lf_minus(up3));
    return hn + hp;
}

// ---------------------------------------------------------------------------------------------
// v1 ("direct"): one thread per (i,j) column marching z; every thread evaluates its own faces
// (x-, x+, y-, y+, z+; z- is carried).  Correct and simple -- the parity anchor for the
// optimised kernel below, and the fallback for shapes that one does not cover.
// ---------------------------------------------------------------------------------------------
template <typename REAL, int W, int MODE, int CHUNK>
__global__ void __launch_bounds__(128)
burgers3d_stage_direct(BurgersArgs<REAL> a, const REAL* __restrict__ u, const REAL* __restrict__ uo,
                       REAL* __restrict__ out, REAL* __restrict__ rowacc)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k0 = a.kb + blockIdx.z * CHUNK;
    const int k1 = min(k0 + CHUNK, a.ke);
    if (i >= a.nx || j >= a.ny || k0 >= k1) return;
    const long long col = (long long)i + a.pitch * j;
    const long long P = a.pitch, Z = a.plane;

    // which directional contributions this column takes
    bool xact, yact, zact, rk_xy;
    if (MODE == BMODE_MGPU) {
        xact = (j >= 3 && j <= a.ny - 4) && (i >= 2 && i <= a.nx - 4);     // Kernels.cu:250,266
        yact = (i >= 3 && i <= a.nx - 4) && (j >= 2 && j <= a.ny - 4);     // Kernels.cu:323,339
        zact = (i >= 3 && i <= a.nx - 4) && (j >= 3 && j <= a.ny - 4);     // Kernels.cu:397
        rk_xy = true;                                                      // Kernels.cu:478
    } else {
        const bool in = (i >= 3 && i <= a.nx - 4) && (j >= 3 && j <= a.ny - 4);
        xact = yact = zact = rk_xy = in;                                   // kernels.cu:518
    }

    auto ld = [&](int k) -> REAL { return (k >= 0 && k < a.nzs) ? u[col + Z * k] : REAL(0); };
    REAL q0 = ld(k0 - 3), q1 = ld(k0 - 2), q2 = ld(k0 - 1), q3 = ld(k0), q4 = ld(k0 + 1), q5 = ld(k0 + 2), q6;
    REAL hz_m = REAL(0);
    bool have_hz = false;

    for (int k = k0; k < k1; k++) {
        q6 = ld(k + 3);
        const long long o = col + Z * k;
        const bool owned = (k >= a.own_lo && k < a.own_hi);
        const bool rk = rk_xy && (MODE == BMODE_MGPU ? true : owned);
        REAL res = q3;
        if (rk) {
            REAL Lu = REAL(0);
            if (owned) {
                if (xact) {
                    const REAL fp = face_flux_ref<REAL, W>(u[o - 2], u[o - 1], q3, u[o + 1], u[o + 2], u[o + 3]);
                    const REAL fm = (MODE == BMODE_MGPU && i == 2) ? REAL(0)
                                  : face_flux_ref<REAL, W>(u[o - 3], u[o - 2], u[o - 1], q3, u[o + 1], u[o + 2]);
                    Lu = -(fp - fm) / a.dx;
                } else if (MODE == BMODE_MGPU && yact && a.stage > 1) {
                    Lu = rowacc[(long long)i + P * k];      // j == 2 row: Lu keeps accumulating within a step
                }
                if (yact) {
                    const REAL gp = face_flux_ref<REAL, W>(u[o - 2 * P], u[o - P], q3, u[o + P], u[o + 2 * P], u[o + 3 * P]);
                    const REAL gm = (MODE == BMODE_MGPU && j == 2) ? REAL(0)
                                  : face_flux_ref<REAL, W>(u[o - 3 * P], u[o - 2 * P], u[o - P], q3, u[o + P], u[o + 2 * P]);
                    Lu -= (gp - gm) / a.dy;
                    if (MODE == BMODE_MGPU && j == 2) rowacc[(long long)i + P * k] = Lu;
                }
                if (zact) {
                    if (!have_hz) hz_m = face_flux_ref<REAL, W>(q0, q1, q2, q3, q4, q5);
                    const REAL hz_p = face_flux_ref<REAL, W>(q1, q2, q3, q4, q5, q6);
                    Lu -= (hz_p - hz_m) / a.dz;
                    hz_m = hz_p; have_hz = true;
                    if (MODE == BMODE_SGPU && a.visc) {
                        const REAL term =
                            a.kx * (-u[o - 2] + 16 * u[o - 1] - 30 * q3 + 16 * u[o + 1] - u[o + 2]) +
                            a.ky * (-u[o - 2 * P] + 16 * u[o - P] - 30 * q3 + 16 * u[o + P] - u[o + 2 * P]) +
                            a.kz * (-q1 + 16 * q2 - 30 * q3 + 16 * q4 - q5);
                        Lu += term;
                        if (k >= a.quirk_lo && k <= a.quirk_hi) Lu += term;
                    }
                }
            } else {
                have_hz = false;
            }
            const REAL qo = (a.stage == 1) ? q3 : uo[o];
            res = rk_combine<REAL>(a.stage, q3, qo, Lu, a.dt);
        }
        out[o] = res;
        q0 = q1; q1 = q2; q2 = q3; q3 = q4; q4 = q5; q5 = q6;
    }
}

}  // namespace adv
