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

// One launch per RK stage on a multi-rank slab (3-D targets): the launch's z chunks are, in this order, up to two
// BOUNDARY windows (the 3 planes next to each neighbour: computed first, stored locally and into the neighbour's
// ghost planes through the NVLink mapping) and then the chunks of the interior window.  A boundary window first
// waits (one thread spins, the CTA follows) until the neighbour has raised this rank's flag to wait_val -- its
// previous stage has landed in our ghost planes and it has stopped reading the ghost planes we are about to
// overwrite -- and the last of its CTAs to finish raises the neighbour's flag to signal_val.  nb == 0: plain launch.
struct SlabSched {
    int nb;                             // boundary windows in this launch: blockIdx.z = 0 .. nb-1
    int wb[2], we[2];                   // their plane ranges [wb, we)
    void* out2[2];                      // second destination of a boundary window (neighbour's arena buffer)
    long long off2[2];                  // element offset of its ghost planes relative to our planes
    const unsigned* wait_flag[2];       // local flag word the window waits on (null: no wait)
    unsigned* peer_flag[2];             // neighbour's flag word the window raises when complete
    unsigned* ticket[2];                // local CTA counters (zero between launches)
    unsigned wait_val, signal_val;
    int ib, ie, chunk;                  // interior window and its chunk length
};

// window of this CTA: returns the boundary index (0/1) or -1 for an interior chunk.  (The two-element arrays are
// read with selects, never with a run-time index: the struct stays in the kernel-parameter constant bank.)
__device__ __forceinline__ int slab_window(const SlabSched& sc, int& kb, int& ke)
{
    const int z = blockIdx.z;
    if (z < sc.nb) { kb = z ? sc.wb[1] : sc.wb[0]; ke = z ? sc.we[1] : sc.we[0]; return z; }
    kb = sc.ib + (z - sc.nb) * sc.chunk;
    ke = min(kb + sc.chunk, sc.ie);
    return -1;
}
__device__ __forceinline__ void* slab_out2(const SlabSched& sc, int bw, void* dflt) { return bw < 0 ? dflt : (bw ? sc.out2[1] : sc.out2[0]); }
__device__ __forceinline__ long long slab_off2(const SlabSched& sc, int bw, long long dflt) { return bw < 0 ? dflt : (bw ? sc.off2[1] : sc.off2[0]); }
__device__ __forceinline__ void slab_wait(const SlabSched& sc, int bw, bool leader)
{
    if (bw < 0) return;
    const volatile unsigned* f = bw ? sc.wait_flag[1] : sc.wait_flag[0];
    if (!f) return;
    if (leader) {
        while ((int)(*f - sc.wait_val) < 0) __nanosleep(64);
        __threadfence();
    }
    __syncthreads();
}
// call with every thread of the CTA, after its last store (the window is taken from blockIdx again so that
// nothing has to stay in a register across the kernel's main loop)
__device__ __forceinline__ void slab_signal(const SlabSched& sc, bool leader)
{
    const int z = blockIdx.z;
    if (z >= sc.nb) return;
    __threadfence_system();             // this thread's peer stores are visible system-wide before the ticket
    __syncthreads();
    if (leader) {
        unsigned* ticket = z ? sc.ticket[1] : sc.ticket[0];
        unsigned* flag = z ? sc.peer_flag[1] : sc.peer_flag[0];
        const unsigned n = gridDim.x * gridDim.y;
        if (atomicAdd(ticket, 1u) == n - 1) {
            *ticket = 0;
            __threadfence_system();
            *(volatile unsigned*)flag = sc.signal_val;
            __threadfence_system();
        }
    }
}

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
