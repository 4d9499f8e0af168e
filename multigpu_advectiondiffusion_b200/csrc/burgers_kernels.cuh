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

// =============================================================================================
// Ring kernel (burgers3d_stage_ring): every WENO reconstruction is evaluated exactly once, and the
// inner loop is shaped around the issue port (an FP64 instruction holds it for two cycles, everything
// else for one: tools/fp64_mix.cu).  Development history and measurements: DESIGN.md section 3.2.
//
//  * Pair formulation.  f_{c+1/2} = hn(c) + hp(c+1): hn(c) (minus side, from f+ on c-2..c+2) and hp(c)
//    (plus side of face c-1/2, from f- on c-2..c+2) are both centred on cell c, and by symmetry
//    hp(a,b,c,d,e) == hn(e,d,c,b,a): one routine (weno_side6) serves both.  A thread evaluates one
//    (hn,hp) pair per direction for each of its cells and gets the neighbours' halves through warp
//    shuffles (x), shared memory (y) or its own shared-memory column (z).  The 2*(32+TY) halves on
//    the tile rim are the job of one extra "edge" warp with its own loop.
//  * (f+,f-) plane ring.  u of tile + halo 3 is fetched one plane ahead, split once
//    (f+- = {3/4,-1/4} u^2 by sign, exact) and parked as double2 in a 7-slot shared-memory ring; one
//    LDS.128 returns both.  Iteration m finishes plane m: planes m-2..m+3 are live, m+4 is staged.
//    The z stencil reads the thread's own column of the ring: only the lower z-face flux and two x
//    partial sums are loop-carried.
//  * WENO arithmetic (weno_side6): 6h = (-b+5c+2d) + [2(t0-t1)n0 + (t1-t2)n2] / (n0+n1+n2),
//    n_k = prod_{m!=k} P_m, P_k = (eps+beta_k)^2/d_k: one reciprocal (rcp.approx + 2 Newton steps)
//    per side instead of the reference's 7 divisions, ~43 FP64 instructions per side; 1/6 and
//    1/dx,1/dy,1/dz are folded into one multiplier.  Same formulas, different rounding: parity is to
//    1e-12, not bitwise (tests/test_gpu_parity.py).
//  * Split-phase barrier.  One mbarrier per plane: a thread arrives after staging plane m+4 and waits
//    only after it has computed the z faces of plane m+1, which depend on nothing other threads wrote
//    (each compute thread stages its own cells; halo points are staged by whoever is free).
//  * Steady state.  Planes whose window/ownership flags are all on, in tiles inside the update region,
//    run a body compiled without any of those tests (FAST); rim tiles and the first/last planes of a
//    chunk run the general body.  RK stage, WENO flavour, target and the peer store are template
//    parameters; 32x16 tile, 2 cells per thread (rows ty, ty+8), 8+1 warps, 2 CTAs/SM.
// =============================================================================================
template <typename REAL> struct V2T;
template <> struct V2T<double> { typedef double2 type; };
template <> struct V2T<float> { typedef float2 type; };

__device__ __forceinline__ double fast_rcp(double den)
{
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(den));
    double e = fma(-den, y, 1.0); y = fma(y, e, y);
    e = fma(-den, y, 1.0); y = fma(y, e, y);
    return y;
}
__device__ __forceinline__ float fast_rcp(float den) { return __frcp_rn(den); }

__device__ __forceinline__ void mbar_init(unsigned addr, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(addr), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned addr)
{
    asm volatile("{\n.reg .b64 st;\nmbarrier.arrive.shared::cta.b64 st, [%0];\n}" ::"r"(addr) : "memory");
}
__device__ __forceinline__ void mbar_wait_hint(unsigned addr, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n"
        "@P1 bra MBAR_DONE;\n"
        "bra MBAR_WAIT;\n"
        "MBAR_DONE:\n"
        "}\n" ::"r"(addr), "r"(parity), "r"(0x989680u) : "memory");
}

template <typename REAL, int TY>
struct BurgersRingSmem {
    static constexpr int TX = 32, H = 3, NSLOT = 7;
    static constexpr int RP = TX + 2 * H, ROWS = TY + 2 * H, NPTS = RP * ROWS;
    static constexpr int NHALO = 2 * H * RP + 2 * H * TY;
    typename V2T<REAL>::type slot[NSLOT][NPTS];
    REAL hn_y[2][TY + 2][TX], hp_y[2][TY + 2][TX];     // [.][r+1][.] <-> tile row r
    REAL edge_hn_x[2][TY], edge_hp_x[2][TY];           // hn of column -1, hp of column TX
    unsigned long long bar;
};

template <typename REAL, int STAGE>
__device__ __forceinline__ REAL rk_stage(REAL u_m, REAL uo_m, REAL Lu, REAL dt, const REAL* __restrict__ wc)
{
    if (STAGE == 1) return u_m + dt * Lu;
    const REAL t = u_m + dt * Lu;
    if (STAGE == 2) return wc[14] * uo_m + wc[15] * t;
    return (uo_m + REAL(2) * t) * wc[19];
}

// SIX TIMES the one-sided WENO5 value: 6h = (-b + 5c + 2d) + [2 (t0-t1) n0 + (t1-t2) n2] / (n0+n1+n2).
// The candidate polynomial then has small-integer coefficients (immediates, no constant loads) and the
// factor 6 of the denominator disappears; the caller multiplies the flux difference by 1/(6 dx).
template <typename REAL, int W>
__device__ __forceinline__ REAL weno_side6(REAL a, REAL b, REAL c, REAL d, REAL e, const REAL* __restrict__ wc)
{
    const REAL t0 = (a - 2 * b) + c, s0 = 3 * c + (a - 4 * b);
    const REAL t1 = (b - 2 * c) + d, s1 = b - d;
    const REAL t2 = (c - 2 * d) + e, s2 = 3 * c + (e - 4 * d);
    const REAL q6 = 2 * d + (5 * c - b);
    const REAL E0 = (wc[0] * t0) * t0 + ((wc[3] * s0) * s0 + wc[6]);
    const REAL E1 = (wc[1] * t1) * t1 + ((wc[4] * s1) * s1 + wc[7]);
    const REAL E2 = (wc[2] * t2) * t2 + ((wc[5] * s2) * s2 + wc[8]);
    const REAL d01 = t0 - t1, d12 = t1 - t2;
    if (W == WENO_JS) {
        const REAL P0 = E0 * E0, P1 = E1 * E1, P2 = E2 * E2;
        const REAL n0 = P1 * P2, n1 = P0 * P2, n2 = P0 * P1;
        const REAL A = d01 * n0;
        const REAL num = d12 * n2 + (A + A);
        return num * fast_rcp((n0 + n1) + n2) + q6;
    } else {
        const REAL tau = fabs(E0 - E2);
        const REAL A0 = (wc[16] * (E0 + tau)) * (E1 * E2);
        const REAL A1 = (wc[17] * (E1 + tau)) * (E0 * E2);
        const REAL A2 = (wc[18] * (E2 + tau)) * (E0 * E1);
        const REAL A = d01 * A0;
        const REAL num = d12 * A2 + (A + A);
        return num * fast_rcp((A0 + A1) + A2) + q6;
    }
}

// reciprocal from the hardware seed (rcp.approx.ftz.f64, ~2^-23) and ONE cubic correction step:
// y = y0 (1 + e + e^2), e = 1 - den*y0  ->  relative error e^3 plus rounding; three FMAs instead of the four of two
// Newton steps
__device__ __forceinline__ double fast_rcp3(double den)
{
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(den));
    const double e = fma(-den, y, 1.0);
    const double r = fma(e, e, e);
    return fma(y, r, y);
}
__device__ __forceinline__ float fast_rcp3(float den) { return __frcp_rn(den); }

// SIX TIMES the one-sided WENO5 value centred on c, from the FORWARD differences along the line
//   g0 = b-a, g1 = c-b, g2 = d-c, g3 = e-d   and   t_k = g_{k+1} - g_k   of the five values (a, b, c, d, e).
// REV = false: the value weno_side6(a, b, c, d, e) stands for; REV = true: weno_side6(e, d, c, b, a) -- the same
// differences read backwards (signs folded into the FMAs), so a cell's minus- and plus-side halves and the halves
// of the next cell along the line share their differences.  Same formulas as weno_side6, 40 FP64 instructions
// instead of 43: differences of differences (14 -> 13), the halved P_0 (see BurgersArgs::wp), one cubic
// reciprocal step.  `we` = BurgersArgs::wp (JS) or ::wc (Z).
template <typename REAL, int W, bool REV>
__device__ __forceinline__ REAL weno_core6(REAL c, REAL g0, REAL g1, REAL g2, REAL g3, REAL t0, REAL t1, REAL t2,
                                           const REAL* __restrict__ we, const REAL* __restrict__ wc)
{
    const REAL T0 = REV ? t2 : t0, T2 = REV ? t0 : t2;
    const REAL sa = fma(REAL(2), g1, t0);       // a - 4b + 3c
    const REAL sb = fma(REAL(-2), g2, t2);      // 3c - 4d + e
    const REAL S0 = REV ? sb : sa, S2 = REV ? sa : sb;
    const REAL S1 = g1 + g2;                    // d - b (enters squared)
    const REAL q6 = REV ? fma(REAL(6), c, -fma(REAL(2), g1, g2)) : fma(REAL(6), c, fma(REAL(2), g2, g1));
    (void)g0; (void)g3;
    const REAL E0 = (we[0] * T0) * T0 + ((we[3] * S0) * S0 + we[6]);
    const REAL E1 = (we[1] * t1) * t1 + ((we[4] * S1) * S1 + we[7]);
    const REAL E2 = (we[2] * T2) * T2 + ((we[5] * S2) * S2 + we[8]);
    const REAL d01 = T0 - t1, d12 = t1 - T2;
    if (W == WENO_JS) {
        const REAL P0 = E0 * E0, P1 = E1 * E1, P2 = E2 * E2;      // P0 is half its textbook value
        const REAL n0 = P1 * P2, n1 = P0 * P2, n2 = P0 * P1;
        const REAL num = fma(d12, n2, d01 * n0);
        const REAL den = fma(REAL(0.5), n0, n1 + n2);
        return fma(num, fast_rcp3(den), q6);
    } else {
        const REAL tau = fabs(E0 - E2);
        const REAL A0 = (wc[16] * (E0 + tau)) * (E1 * E2);
        const REAL A1 = (wc[17] * (E1 + tau)) * (E0 * E2);
        const REAL A2 = (wc[18] * (E2 + tau)) * (E0 * E1);
        const REAL A = d01 * A0;
        const REAL num = fma(d12, A2, A + A);
        return fma(num, fast_rcp3((A0 + A1) + A2), q6);
    }
}

template <typename REAL> __device__ __forceinline__ bool is_neg(REAL v);
template <> __device__ __forceinline__ bool is_neg<double>(double v) { return __double2hiint(v) < 0; }
template <> __device__ __forceinline__ bool is_neg<float>(float v) { return __float_as_int(v) < 0; }

template <bool B> struct BoolTag { static constexpr bool value = B; };

template <typename REAL, int W, int MODE, int TY, int CPT, int MINB, int STAGE, bool PEER>
__global__ void __launch_bounds__(32 * (TY / CPT + 1), MINB)
burgers3d_stage_ring(BurgersArgs<REAL> a, int chunk, const REAL* __restrict__ u, const REAL* __restrict__ uo,
                      REAL* __restrict__ out, REAL* __restrict__ rowacc)
{
    typedef typename V2T<REAL>::type V2;
    typedef BurgersRingSmem<REAL, TY> S;
    constexpr int TX = S::TX, H = S::H, RP = S::RP, NSLOT = S::NSLOT, NHALO = S::NHALO;
    constexpr int NW = TY / CPT;
    constexpr int NT = 32 * (NW + 1);
    constexpr int NPH = (NHALO + NT - 1) / NT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    S& sm = *reinterpret_cast<S*>(smem_raw);

    const int tx = threadIdx.x, ty = threadIdx.y;
    const int tid = ty * 32 + tx;
    const int i0 = blockIdx.x * TX, j0 = blockIdx.y * TY;
    const int kb = a.kb + blockIdx.z * chunk;
    const int ke = min(kb + chunk, a.ke);
    if (kb >= ke) return;
    const long long P = a.pitch, Z = a.plane;
    const REAL* __restrict__ wc = a.wc;
    const unsigned bar = (unsigned)__cvta_generic_to_shared(&sm.bar);
    const bool interior = (i0 >= 3) && (i0 + TX - 1 <= a.nx - 4) && (j0 >= 3) && (j0 + TY - 1 <= a.ny - 4);
    // steady-state planes: finish m, own m and m+1, stage m+4, read u(m+1) -- all inside the chunk and the slab
    const int f0 = max(kb, a.own_lo), f1 = min(min(ke - 2, a.own_hi - 2), a.nzs - 5);

    // halo points of the tile staged by this thread
    int h_idx[NPH];
    long long h_off[NPH];
    bool h_ok[NPH];
#pragma unroll
    for (int n = 0; n < NPH; n++) {
        const int hh = tid + n * NT;
        int r, cc;
        if (hh < H * RP) { r = hh / RP; cc = hh - r * RP; }
        else if (hh < 2 * H * RP) { const int t = hh - H * RP; r = t / RP; cc = t - r * RP; r += H + TY; }
        else { const int t = hh - 2 * H * RP; r = H + t / (2 * H); const int q = t - (t / (2 * H)) * (2 * H); cc = q < H ? q : TX + q; }
        const int gi = i0 - H + cc, gj = j0 - H + r;
        h_idx[n] = (hh < NHALO) ? r * RP + cc : -1;
        h_ok[n] = hh < NHALO && gi >= 0 && gi < a.nx && gj >= 0 && gj < a.ny;
        h_off[n] = h_ok[n] ? (long long)gi + P * gj : 0;
    }
    auto split = [&](REAL v) -> V2 {
        const REAL sq = v * v;
        const bool ng = is_neg<REAL>(v);
        V2 r;
        r.x = (ng ? -wc[15] : wc[14]) * sq;
        r.y = (ng ? wc[14] : -wc[15]) * sq;
        return r;
    };
    REAL st_halo[NPH];
    auto fetch_halo = [&](auto fast, int p) {
        constexpr bool FAST = decltype(fast)::value;
        const REAL* __restrict__ up = u + Z * p;
        const bool pl = FAST || (p >= 0 && p < a.nzs);
#pragma unroll
        for (int n = 0; n < NPH; n++) {
            if (FAST) { if ((n + 1) * NT <= NHALO || h_idx[n] >= 0) st_halo[n] = up[h_off[n]]; }
            else st_halo[n] = (pl && h_ok[n]) ? up[h_off[n]] : REAL(0);
        }
    };
    auto park_halo = [&](int sidx) {
        V2* __restrict__ sp = sm.slot[sidx];
#pragma unroll
        for (int n = 0; n < NPH; n++) if ((n + 1) * NT <= NHALO || h_idx[n] >= 0) sp[h_idx[n]] = split(st_halo[n]);
    };
    auto slot_of = [&](int s0, int d) -> int { int t = s0 + d; return t >= NSLOT ? t - NSLOT : t; };

    if (tid == 0) mbar_init(bar, NT);

    if (ty == NW) {
        // =============================== edge warp ===============================
        for (int d = 0; d < 6; d++) { fetch_halo(BoolTag<false>(), kb - 3 + d); park_halo(d); }
        __syncthreads();
        mbar_arrive(bar);
        unsigned parity = 0;
        int s0 = 0;
        for (int m = kb - 1; m < ke; m++) {
            const bool do_pairs = (m + 1 < ke) && (m + 1 >= a.own_lo && m + 1 < a.own_hi);
            const bool do_fetch = (m + 4 <= ke + 2);
            const int nb = (m & 1) ^ 1;
            if (do_fetch) fetch_halo(BoolTag<false>(), m + 4);
            mbar_wait_hint(bar, parity);
            parity ^= 1u;
            if (do_pairs) {
                const V2* __restrict__ s3 = sm.slot[slot_of(s0, 3)];
#pragma unroll 1
                for (int it = tx; it < 2 * (TX + TY); it += 32) {
                    int oc, st;
                    bool minus;
                    if (it < TY)               { oc = (it + H) * RP + (H - 1);                st = 1;  minus = true;  }
                    else if (it < 2 * TY)      { oc = (it - TY + H) * RP + (H + TX);          st = 1;  minus = false; }
                    else if (it < 2 * TY + TX) { oc = (H - 1) * RP + (it - 2 * TY + H);       st = RP; minus = true;  }
                    else                       { oc = (H + TY) * RP + (it - 2 * TY - TX + H); st = RP; minus = false; }
                    const V2 m2 = s3[oc - 2 * st], m1 = s3[oc - st], cc = s3[oc], p1 = s3[oc + st], p2 = s3[oc + 2 * st];
                    const REAL v = weno_side6<REAL, W>(minus ? m2.x : p2.y, minus ? m1.x : p1.y, minus ? cc.x : cc.y,
                                                        minus ? p1.x : m1.y, minus ? p2.x : m2.y, wc);
                    if (it < TY) sm.edge_hn_x[nb][it] = v;
                    else if (it < 2 * TY) sm.edge_hp_x[nb][it - TY] = v;
                    else if (it < 2 * TY + TX) sm.hn_y[nb][0][it - 2 * TY] = v;
                    else sm.hp_y[nb][TY + 1][it - 2 * TY - TX] = v;
                }
            }
            if (do_fetch) park_halo(slot_of(s0, 6));
            mbar_arrive(bar);
            s0 = (s0 + 1 == NSLOT) ? 0 : s0 + 1;
        }
        return;
    }

    // =============================== compute warps ===============================
    const int i = i0 + tx;
    bool in_dom[CPT];
    long long col0 = (long long)i + P * (j0 + ty);     // cell c lives at col0 + c*NW*P
#pragma unroll
    for (int c = 0; c < CPT; c++) in_dom[c] = i < a.nx && (j0 + ty + c * NW) < a.ny;
    if (!in_dom[0]) col0 = 0;
    const int own0 = (ty + H) * RP + tx + H;
    constexpr int CS = NW * RP;                        // slot offset between a thread's cells
    // per-cell activity of the general body, packed once into a bit mask (4 bits per cell: x, y, z
    // contribution active, j==2 scratch row) -- recomputing the index comparisons every plane costs
    // issue slots the rim tiles do not have either
    unsigned act = 0;
    {
        const bool i_in3 = (i >= 3 && i <= a.nx - 4);
#pragma unroll
        for (int c = 0; c < CPT; c++) {
            const int j = j0 + ty + c * NW;
            const bool j_in3 = (j >= 3 && j <= a.ny - 4);
            const bool xa = (MODE == BMODE_MGPU) ? (j_in3 && i >= 2 && i <= a.nx - 4) : (i_in3 && j_in3);
            const bool ya = (MODE == BMODE_MGPU) ? (i_in3 && j >= 2 && j <= a.ny - 4) : (i_in3 && j_in3);
            const bool za = i_in3 && j_in3;
            const bool row2 = (MODE == BMODE_MGPU) && j == 2 && ya;
            act |= ((xa ? 1u : 0u) | (ya ? 2u : 0u) | (za ? 4u : 0u) | (row2 ? 8u : 0u)) << (4 * c);
        }
    }
    const bool x_seed0 = (MODE == BMODE_MGPU) && i == 2;

    REAL st_own[CPT];
    auto fetch_own = [&](auto fast, int p) {
        constexpr bool FAST = decltype(fast)::value;
        const REAL* __restrict__ up = u + Z * p + col0;
        const bool pl = FAST || (p >= 0 && p < a.nzs);
#pragma unroll
        for (int c = 0; c < CPT; c++) st_own[c] = (FAST || (pl && in_dom[c])) ? up[(long long)c * NW * P] : REAL(0);
    };
    auto park_own = [&](int sidx) {
        V2* __restrict__ sp = sm.slot[sidx];
#pragma unroll
        for (int c = 0; c < CPT; c++) sp[own0 + c * CS] = split(st_own[c]);
    };
    auto zface = [&](const V2* p0, const V2* p1, const V2* p2, const V2* p3, const V2* p4, const V2* p5, int own) -> REAL {
        const V2 v0 = p0[own], v1 = p1[own], v2 = p2[own], v3 = p3[own], v4 = p4[own], v5 = p5[own];
        return weno_side6<REAL, W>(v0.x, v1.x, v2.x, v3.x, v4.x, wc) + weno_side6<REAL, W>(v5.y, v4.y, v3.y, v2.y, v1.y, wc);
    };
    auto visc13 = [&](const REAL* __restrict__ pc, REAL c0) -> REAL {
        return a.kx * (-pc[-2] + 16 * pc[-1] - 30 * c0 + 16 * pc[1] - pc[2]) +
               a.ky * (-pc[-2 * P] + 16 * pc[-P] - 30 * c0 + 16 * pc[P] - pc[2 * P]) +
               a.kz * (-pc[-2 * Z] + 16 * pc[-Z] - 30 * c0 + 16 * pc[Z] - pc[2 * Z]);
    };

    for (int d = 0; d < 6; d++) {
        fetch_own(BoolTag<false>(), kb - 3 + d); fetch_halo(BoolTag<false>(), kb - 3 + d);
        park_own(d); park_halo(d);
    }
    __syncthreads();
    REAL Fz_prev[CPT], X1[CPT], X2[CPT], u_nxt[CPT], uo_nxt[CPT];
#pragma unroll
    for (int c = 0; c < CPT; c++) {
        u_nxt[c] = uo_nxt[c] = X1[c] = X2[c] = REAL(0);
        Fz_prev[c] = zface(sm.slot[0], sm.slot[1], sm.slot[2], sm.slot[3], sm.slot[4], sm.slot[5], own0 + c * CS);
    }
    long long om = col0 + Z * (kb - 2);                // element offset of cell 0 in plane m, bumped per iteration
    mbar_arrive(bar);
    unsigned parity = 0;
    int s0 = 0;

    auto iteration = [&](auto fast, int m) {
        constexpr bool FAST = decltype(fast)::value;
        const bool do_final = FAST || (m >= kb);
        const bool own_m = FAST || (m >= a.own_lo && m < a.own_hi);
        const bool do_z = FAST || (do_final && (m >= a.own_lo - 1 && m < a.own_hi));
        const bool do_pairs = FAST || ((m + 1 < ke) && (m + 1 >= a.own_lo && m + 1 < a.own_hi));
        const bool do_fetch = FAST || (m + 4 <= ke + 2);
        const int eb = m & 1, nb = eb ^ 1;
        om += Z;
        if (do_fetch) { fetch_own(fast, m + 4); fetch_halo(fast, m + 4); }
        REAL u_m[CPT], uo_m[CPT];
#pragma unroll
        for (int c = 0; c < CPT; c++) {
            u_m[c] = u_nxt[c]; uo_m[c] = uo_nxt[c];
            if (FAST || (in_dom[c] && m + 1 < ke)) {
                u_nxt[c] = u[om + Z + (long long)c * NW * P];
                if (STAGE > 1) uo_nxt[c] = uo[om + Z + (long long)c * NW * P];
            }
        }
        REAL Fz[CPT];
#pragma unroll
        for (int c = 0; c < CPT; c++) Fz[c] = REAL(0);
        if (do_z) {
            const V2 *p0 = sm.slot[slot_of(s0, 0)], *p1 = sm.slot[slot_of(s0, 1)], *p2 = sm.slot[slot_of(s0, 2)],
                     *p3 = sm.slot[slot_of(s0, 3)], *p4 = sm.slot[slot_of(s0, 4)], *p5 = sm.slot[slot_of(s0, 5)];
#pragma unroll
            for (int c = 0; c < CPT; c++) Fz[c] = zface(p0, p1, p2, p3, p4, p5, own0 + c * CS);
        }
        mbar_wait_hint(bar, parity);
        parity ^= 1u;

        if (do_final) {
#pragma unroll
            for (int c = 0; c < CPT; c++) {
                const int r = ty + c * NW;
                const long long o = om + (long long)c * NW * P;
                REAL Lu = REAL(0);
                if (FAST || interior) {
                    if (own_m) {
                        const REAL fu = (tx == 31) ? X1[c] + sm.edge_hp_x[eb][r] : X1[c];
                        const REAL fu_old = (tx == 0) ? sm.edge_hn_x[eb][r] + X2[c] : X2[c];
                        const REAL gu = sm.hn_y[eb][r + 1][tx] + sm.hp_y[eb][r + 2][tx];
                        const REAL gu_old = sm.hn_y[eb][r][tx] + sm.hp_y[eb][r + 1][tx];
                        Lu = -(fu - fu_old) * a.rdx6;
                        Lu -= (gu - gu_old) * a.rdy6;
                        Lu -= (Fz[c] - Fz_prev[c]) * a.rdz6;
                        if (MODE == BMODE_SGPU && a.visc) {
                            const REAL term = visc13(u + o, u_m[c]);
                            Lu += term;
                            if (m >= a.quirk_lo && m <= a.quirk_hi) Lu += term;
                        }
                    }
                    const bool rk = (MODE == BMODE_MGPU) ? true : own_m;
                    const REAL res = rk_stage<REAL, STAGE>(u_m[c], uo_m[c], Lu, a.dt, wc);
                    const REAL val = rk ? res : u_m[c];
                    out[o] = val;
                    if (PEER) a.out2[o + a.off2] = val;
                } else {
                    const bool xa = (act >> (4 * c)) & 1u, ya = (act >> (4 * c)) & 2u, za = (act >> (4 * c)) & 4u;
                    const bool row2 = (act >> (4 * c)) & 8u;
                    if (own_m) {
                        const REAL fu = (tx == 31) ? X1[c] + sm.edge_hp_x[eb][r] : X1[c];
                        REAL fu_old = (tx == 0) ? sm.edge_hn_x[eb][r] + X2[c] : X2[c];
                        fu_old = x_seed0 ? REAL(0) : fu_old;
                        const REAL gu = sm.hn_y[eb][r + 1][tx] + sm.hp_y[eb][r + 2][tx];
                        REAL gu_old = sm.hn_y[eb][r][tx] + sm.hp_y[eb][r + 1][tx];
                        gu_old = ((MODE == BMODE_MGPU) && (j0 + r == 2)) ? REAL(0) : gu_old;
                        Lu = xa ? -(fu - fu_old) * a.rdx6 : REAL(0);
                        Lu = ya ? Lu - (gu - gu_old) * a.rdy6 : Lu;
                    }
                    if (in_dom[c]) {
                        const bool rk = (MODE == BMODE_MGPU) ? true : (own_m && za);
                        if (own_m) {
                            if (row2) {
                                const long long ra = (long long)i + P * m;
                                if (STAGE > 1) Lu += rowacc[ra];
                                rowacc[ra] = Lu;
                            }
                            Lu = za ? Lu - (Fz[c] - Fz_prev[c]) * a.rdz6 : Lu;
                            if (MODE == BMODE_SGPU && a.visc && za) {
                                const REAL term = visc13(u + o, u_m[c]);
                                Lu += term;
                                if (m >= a.quirk_lo && m <= a.quirk_hi) Lu += term;
                            }
                        }
                        const REAL res = rk_stage<REAL, STAGE>(u_m[c], uo_m[c], Lu, a.dt, wc);
                        const REAL val = rk ? res : u_m[c];
                        out[o] = val;
                        if (PEER) a.out2[o + a.off2] = val;
                    }
                }
                Fz_prev[c] = Fz[c];
            }
        }
        if (do_pairs) {
            const V2* __restrict__ s3 = sm.slot[slot_of(s0, 3)];
#pragma unroll
            for (int c = 0; c < CPT; c++) {
                const int r = ty + c * NW;
                const int own = own0 + c * CS;
                const V2 cc = s3[own];
                const V2 xm2 = s3[own - 2], xm1 = s3[own - 1], xp1 = s3[own + 1], xp2 = s3[own + 2];
                const REAL hn_x = weno_side6<REAL, W>(xm2.x, xm1.x, cc.x, xp1.x, xp2.x, wc);
                const REAL hp_x = weno_side6<REAL, W>(xp2.y, xp1.y, cc.y, xm1.y, xm2.y, wc);
                const V2 ym2 = s3[own - 2 * RP], ym1 = s3[own - RP], yp1 = s3[own + RP], yp2 = s3[own + 2 * RP];
                sm.hn_y[nb][r + 1][tx] = weno_side6<REAL, W>(ym2.x, ym1.x, cc.x, yp1.x, yp2.x, wc);
                sm.hp_y[nb][r + 1][tx] = weno_side6<REAL, W>(yp2.y, yp1.y, cc.y, ym1.y, ym2.y, wc);
                const REAL hpR = __shfl_down_sync(0xffffffffu, hp_x, 1);
                const REAL hnL = __shfl_up_sync(0xffffffffu, hn_x, 1);
                X1[c] = (tx == 31) ? hn_x : hn_x + hpR;
                X2[c] = (tx == 0) ? hp_x : hnL + hp_x;
            }
        }
        if (do_fetch) { const int t = slot_of(s0, 6); park_own(t); park_halo(t); }
        mbar_arrive(bar);
        s0 = (s0 + 1 == NSLOT) ? 0 : s0 + 1;
    };

    for (int m = kb - 1; m < ke; m++) {
        if (interior && m >= f0 && m <= f1) iteration(BoolTag<true>(), m);
        else iteration(BoolTag<false>(), m);
    }
}

}  // namespace adv
