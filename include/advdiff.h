/*
 * advdiff.h -- C ABI of libadvdiff.so: B200 (sm_100a) explicit finite-difference
 * time stepping for the 2-D/3-D heat equation (4th-order Laplacian) and the 3-D
 * Burgers equation (WENO5 + Lax-Friedrichs splitting) inside SSP-RK3.
 *
 * This is the drop-in boundary for the hot path of
 * cfd-learner/MultiGPU_AdvectionDiffusion.  The reference's link surface is a
 * set of `extern "C"` launch wrappers that take dim3 by value and operate on a
 * materialised right-hand-side array Lu (one kernel per sweep, then Compute_RK):
 *
 *   SingleGPU/Diffusion3d_baselineCode/acoustics3d.h:55-60  CopyToConstantMemory, Call_Lu3d, Call_RK3d
 *   SingleGPU/Diffusion2d/diffusion2d.h:55-60               CopyToConstantMemory, Call_Lu2d, Call_RK2d
 *   SingleGPU/Burgers3d_WENO5/Burgers.h:70-83               Call_Adv_x/y/z, Call_Diff_, Call_sspRK
 *   MultiGPU/Diffusion3d_Baseline/DiffusionMPICUDA.h:105-127 Call_Diff_, Call_sspRK, Copy*Async
 *   MultiGPU/Burgers3d_Baseline/BurgersMPI.h:105-133        Compute_Adv_{x,y,z}_Async, Compute_sspRK3Async, Copy*Async
 *
 * Here every RK stage is ONE fused kernel (RHS + stage combine, no Lu array),
 * so the per-sweep wrappers have no counterpart; what replaces them is
 *   (1) stage-level entry points on raw device pointers (adv_*_stage_*), the
 *       direct replacement for one "Call_Lu.. / Call_Adv_.. then Call_RK.." group, and
 *   (2) a solver handle (adv_create ... adv_destroy) that owns the three field
 *       buffers, streams and halo exchange and reproduces each reference
 *       driver's time loop (main.cpp / main.c `while (t < tEnd)`).
 * Plain C: no dim3, no C++ types, no torch types.  All functions return 0 on
 * success or an ADV_ERR_* code; adv_last_error() gives the message (the
 * reference's checkCuda/MPI_CHECK print and exit(-1); callers that want that
 * behaviour print adv_last_error() and exit).
 *
 * Host arrays use the reference's layout: x fastest, o = i + nx*j + nx*ny*k,
 * multi-GPU targets z-padded by RADIUS=3 planes at both ends (nz+6 planes).
 */
#ifndef ADVDIFF_H
#define ADVDIFF_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ADV_VERSION 1
#define ADV_RADIUS 3            /* ghost planes per side; BurgersMPI.h:50 */

/* error codes */
enum {
    ADV_OK = 0,
    ADV_ERR_INVALID = 1,        /* bad argument / unsupported size */
    ADV_ERR_CUDA = 2,           /* CUDA runtime or driver error */
    ADV_ERR_NO_DEVICE = 3,      /* no usable sm_100 device */
    ADV_ERR_COMM = 4,           /* NCCL error / NCCL not loadable */
    ADV_ERR_NOMEM = 5,
    ADV_ERR_STATE = 6           /* call order (e.g. step before upload) */
};

/* which reference program's semantics the solver reproduces */
enum {
    /* SingleGPU/Diffusion2d: 9-pt O4 Laplacian, frame of 3 frozen, dt factor 0.9 */
    ADV_TARGET_HEAT2D_SGPU = 0,
    /* SingleGPU/Diffusion3d_baselineCode: 13-pt O4 on [2,N-3]^3, frame of 2 frozen */
    ADV_TARGET_HEAT3D_SGPU_BASELINE = 1,
    /* SingleGPU/Diffusion3d_Blocking: Laplacian on x,y in [3,N-4], z in [3,nz-3];
     * RK on [2,N-3]^3 with zero RHS elsewhere */
    ADV_TARGET_HEAT3D_SGPU_BLOCKING = 2,
    /* MultiGPU/Diffusion3d_Baseline: z-padded, kx/12-isotropic operator (Kernels.cu:240-243),
     * dt factor 0.8, RK on x,y in [3,N-4] for every stored plane */
    ADV_TARGET_HEAT3D_MGPU = 3,
    /* SingleGPU/Burgers3d_WENO5: WENO5-Z + C=1e-5 viscous term, frame of 3 frozen.  Also accepted with
     * nranks > 1 (no counterpart in the reference, whose multi-GPU Burgers program is the JS/inviscid one below):
     * the same semantics on z-slabs -- the frozen frame planes 0..2 and nz-3..nz-1 stand in for pad planes, rank r
     * owns global planes [3 + r*m, 3 + (r+1)*m), m = (nz-6)/nranks, and stores m + 6 planes; results equal the
     * one-rank run bit for bit */
    ADV_TARGET_BURGERS3D_SGPU = 4,
    /* MultiGPU/Burgers3d_Baseline: WENO5-JS inviscid, z-padded, RK on every cell, the
     * i=2 / j=2 pseudo-updates of the reference at np>=2 */
    ADV_TARGET_BURGERS3D_MGPU = 5,
    /* MultiGPU/Diffusion2d_Baseline: y-slabs, y-padded by 3 rows at both ends (ny+6 rows), anisotropic
     * 9-pt operator on columns [3,nx-4] and rows [2, ny+3) (Compute_Laplace's jstart-1 offset,
     * Kernels.cu:144), dt factor 0.8, RK on every cell */
    ADV_TARGET_HEAT2D_MGPU = 6,
    /* MultiGPU/Burgers2d_Baseline: y-slabs, WENO5-JS inviscid, x sweep on cells [2,nx-4] with the flux at
     * 2-1/2 taken as 0 (Kernels.cu:249), y sweep on columns [3,nx-4], RK on every cell */
    ADV_TARGET_BURGERS2D_MGPU = 7
};

enum { ADV_F64 = 0, ADV_F32 = 1 };          /* the reference's USE_FLOAT switch */

/* flags */
#define ADV_FLAG_VISC_ONCE   0x1   /* BURGERS3D_SGPU: add the viscous term once per plane instead of
                                      reproducing the double count on planes 16..18 of
                                      Compute_Laplace_Async (kernels.cu:464-490) */

typedef struct adv_desc {
    int    target;          /* ADV_TARGET_* */
    int    precision;       /* ADV_F64 / ADV_F32 */
    int    nx, ny, nz;      /* PHYSICAL cells (argv Nx Ny Nz); nz ignored for 2-D */
    double lx, ly, lz;      /* argv L W H (single-GPU heat drivers truncate them with atoi: pass the
                               truncated value) */
    double coef;            /* heat: C / K (argv[1]);  Burgers: CFL */
    int    flags;           /* ADV_FLAG_* */
    int    device;          /* CUDA device ordinal of this rank */
    /* z-slab decomposition (MGPU targets only; 1/0 otherwise).  Rank r owns physical planes
     * [r*nz/nranks, (r+1)*nz/nranks) and stores nz/nranks + 6 planes (main.c:64-66).  The two 2-D MGPU
     * targets split y the same way: rows [r*ny/nranks, ...), ny/nranks + 6 stored rows. */
    int    nranks, rank;
    /* 128-byte NCCL unique id shared by all ranks (adv_comm_unique_id), ignored when nranks==1 */
    unsigned char comm_id[128];
} adv_desc;

typedef struct adv_solver adv_solver;

/* ---- library / device ------------------------------------------------------------------ */
const char* adv_last_error(void);
int  adv_version(void);
/* number of CUDA devices with compute capability 10.x (AssignDevices/DeviceScan, Util.cu:43-74) */
int  adv_device_count(int* count);
/* fills a 128-byte NCCL unique id (rank 0 calls it, then distributes the bytes) */
int  adv_comm_unique_id(unsigned char id[128]);

/* ---- solver handle --------------------------------------------------------------------- */
/* replaces: driver set-up, main.c:60-128 / main.cpp:52-86 (constants, cudaMalloc[Pitch] x3,
 * streams, halo buffers).  Collective over ranks when nranks > 1. */
int  adv_create(const adv_desc* desc, adv_solver** out);
int  adv_destroy(adv_solver* s);

/* element counts: the rank-local host array (nx*ny*(nz/nranks+6) for 3-D MGPU, nx*(ny/nranks+6) for 2-D
 * MGPU, nx*ny*nz otherwise) */
size_t adv_local_elems(const adv_solver* s);
/* fixed time step the reference driver would use (main.c:193 / main.cpp:58,130-133) */
double adv_dt(const adv_solver* s);

/* Host -> device of this rank's sub-domain (what Init_subdomain produced, Tools.c:181-200), REAL
 * = double or float per desc.precision.  replaces cudaMemcpy2D H2D, main.c:138 / main.cpp:82. */
int  adv_upload(adv_solver* s, const void* host_local);
/* Device -> host of the rank-local array (all stored planes).  replaces main.c:333. */
int  adv_download(adv_solver* s, void* host_local);
/* Reference initial condition generated on the host with libm exp exactly as the reference
 * does (Call_Init3D / Init_domain case per target), sliced to this rank and uploaded. */
int  adv_init_reference_ic(adv_solver* s);
/* The same closed forms evaluated by a kernel straight into the slab (the reference fills the
 * whole padded domain serially on every rank, main.c:70-73 + Tools.c:124-176).  The device exp() is not the
 * host libm's: values agree with adv_init_reference_ic to the last bit or two, not bit for bit, so parity
 * runs keep the host generator and throughput runs use this one. */
int  adv_init_reference_ic_device(adv_solver* s);

/* nsteps fixed-dt time steps (heat drivers' `for it < iters`).  Asynchronous. */
int  adv_step(adv_solver* s, int nsteps);
/* the drivers' `while (t < tEnd)` loop with the clipped last step (main.c:190-196,
 * main.cpp:127-136); for ADV_TARGET_HEAT3D_MGPU tEnd = dt*iters (main.c:68).  max_steps < 0:
 * unbounded.  Returns the last dt, the step count and the reached time. */
int  adv_run_until(adv_solver* s, double tEnd, int max_steps, double* dt_last, double* steps, double* t_reached);
/* SURVEY 8f-2, no counterpart in the reference's CUDA programs: the same loop with the time step of its Matlab
 * prototype, dt = CFL*dx/max|u| recomputed before every step (LFWENO5FDM3d.m:71) instead of the drivers'
 * max|u| = 1 (main.c:193); max over every stored cell of every rank.  3-D Burgers targets: everything on the
 * device (max|u| folded into stage 3, one 8-byte all-reduce across ranks, dt computed by a kernel, no host
 * synchronisation in the loop), any nranks -- call it on every rank; all ranks return the same dt, step count and
 * time.  2-D Burgers target: one reduction pass + one 8-byte device-to-host read per step, nranks == 1. */
int  adv_run_until_cfl(adv_solver* s, double tEnd, int max_steps, double* dt_last, double* steps, double* t_reached);
int  adv_sync(adv_solver* s);
/* device time (ms) spent in the last adv_step/adv_run_until call, measured with CUDA events on
 * the solver's compute stream; valid after adv_sync */
int  adv_last_elapsed_ms(adv_solver* s, float* ms);
/* number of kernels this library launched on behalf of the handle so far */
long long adv_kernel_launches(const adv_solver* s);
/* raw device pointer of the current solution buffer and its row pitch in elements */
int  adv_device_field(adv_solver* s, void** dptr, size_t* pitch_elems);

/* ---- fused halo exchange over NVLink peer memory (optional; default is NCCL send/recv) ------------ */
/* replaces the reference's pack kernel + D2H + MPI_Isend/Irecv + H2D + unpack kernel per stage
 * (MultiGPU/Burgers3d_Baseline/main.c:208-302, Kernels.cu:40-105): once connected, the boundary-slab stage
 * kernels store their three planes straight into the neighbour GPU's ghost planes and a per-neighbour flag
 * orders the stages; no communication kernel, no host involvement.  Results are bit-identical to the NCCL
 * path.  Call on every rank, after adv_create and before stepping.
 *   adv_peer_handle        64-byte CUDA IPC handle of this rank's device arena (ranks = processes)
 *   adv_peer_connect       map the neighbours' arenas from their handles (NULL where there is no neighbour)
 *   adv_peer_connect_local same for ranks that live in one process (plain-C multi-GPU drivers)
 *   adv_halo_mode          0 single rank, 1 NCCL, 2 fused peer stores */
int  adv_peer_handle(adv_solver* s, unsigned char handle[64]);
int  adv_peer_connect(adv_solver* s, const unsigned char* left_handle, const unsigned char* right_handle);
int  adv_peer_connect_local(adv_solver* s, adv_solver* left, adv_solver* right);
int  adv_halo_mode(const adv_solver* s);
/* The exchange of one stage alone, timed on the device (SURVEY 8d "Halo metric"; the reference's exchange is
 * main.c:208-316): `reps` times, every rank ships its 3 boundary planes per neighbour into the neighbours'
 * ghost planes (the values already there -- the field is unchanged) and waits for theirs.  Collective: call
 * on every rank.  ms_per_exchange = CUDA-event time / reps; bytes_sent_per_exchange = 3 planes x neighbours. */
int  adv_halo_probe(adv_solver* s, int reps, float* ms_per_exchange, double* bytes_sent_per_exchange);

/* One-call convenience used by the drivers: create + upload + run + download + destroy on host
 * buffers (whole padded domain, single rank).  Mirrors one run of a reference binary. */
int  adv_solve_host(const adv_desc* desc, const void* host_in, void* host_out,
                    double tEnd_or_iters, double* dt_last, double* steps, double* t_reached);

/* ---- stage-level entry points on raw device pointers ----------------------------------- */
/* One RK stage = RHS + combine fused, every cell of [0,nx) x [0,ny) x [k_begin,k_end) of u_out
 * is written (cells outside the target's update region are copied from u_in).
 *   stage 1: u_out = uo + dt*L(u_in)                       (uo == u_in is allowed and expected)
 *   stage 2: u_out = 0.75*uo + 0.25*(u_in + dt*L(u_in))
 *   stage 3: u_out = (uo + 2*(u_in + dt*L(u_in)))/3
 * u_out must not alias u_in or uo.  pitch = row stride in elements (>= nx). */
typedef struct adv_stage_args {
    int    target;          /* ADV_TARGET_* (selects operator, frame and quirks) */
    int    flags;
    int    stage;           /* 1, 2, 3 */
    int    nx, ny, nzs;     /* stored extent (nzs includes the 6 pad/ghost planes for MGPU) */
    size_t pitch;
    int    k_begin, k_end;  /* planes of u_out to produce */
    int    own_lo, own_hi;  /* MGPU: planes [own_lo, own_hi) get a right-hand side (3, nzs-3);
                               planes outside are pad planes of the global ends (RK with zero RHS)
                               -- ghost planes that a neighbour rank fills must be excluded from
                               [k_begin,k_end) by the caller */
    int    k_global0;       /* global index of local plane 0 (BURGERS3D_SGPU quirk planes) */
    double dt;
    double dx, dy, dz;      /* Burgers */
    double kx, ky, kz;      /* heat: C/(12 d^2);  Burgers SGPU: viscous C/(12 d^2) */
} adv_stage_args;

/* replaces Call_Lu3d + Call_RK3d (acoustics3d.h:57-59) and Call_Diff_ + Call_sspRK
 * (DiffusionMPICUDA.h:116-121) */
int adv_heat3d_stage_f64(const adv_stage_args* a, const double* u_in, const double* uo, double* u_out, void* cuda_stream);
int adv_heat3d_stage_f32(const adv_stage_args* a, const float* u_in, const float* uo, float* u_out, void* cuda_stream);
/* replaces Call_Lu2d + Call_RK2d (diffusion2d.h:57-59); ny rows, nzs ignored.  With
 * ADV_TARGET_HEAT2D_MGPU it replaces Call_Diff_ + Call_sspRK of MultiGPU/Diffusion2d_Baseline
 * (DiffusionMPICUDA.h:113-118): ny = STORED rows of the slab, [k_begin,k_end) = rows to produce,
 * [own_lo,own_hi) = owned rows, k_global0 = global index of local row 0. */
int adv_heat2d_stage_f64(const adv_stage_args* a, const double* u_in, const double* uo, double* u_out, void* cuda_stream);
int adv_heat2d_stage_f32(const adv_stage_args* a, const float* u_in, const float* uo, float* u_out, void* cuda_stream);
/* replaces Call_Adv_x/y/z + Call_Diff_ + Call_sspRK (Burgers.h:73-82) and
 * Compute_Adv_{x,y,z}_Async + Compute_sspRK3Async (BurgersMPI.h:118-129).
 * rowacc: device scratch of pitch*nzs elements, needed for ADV_TARGET_BURGERS3D_MGPU only: it
 * carries the reference's `Lu -=` accumulation on the j=2 row across the three stages of a step
 * (Kernels.cu:349 with the once-per-step memset of main.c:202).  Stage 1 overwrites it, stages
 * 2 and 3 read and update it; it may be NULL for the other targets. */
int adv_burgers3d_stage_f64(const adv_stage_args* a, const double* u_in, const double* uo, double* u_out, double* rowacc, void* cuda_stream);
int adv_burgers3d_stage_f32(const adv_stage_args* a, const float* u_in, const float* uo, float* u_out, float* rowacc, void* cuda_stream);

/* replaces Compute_Adv_x_Async + Compute_Adv_y_Async + Compute_sspRK3Async of MultiGPU/Burgers2d_Baseline
 * (BurgersMPICUDA.h:118-127), target ADV_TARGET_BURGERS2D_MGPU: ny = STORED rows of the slab,
 * [k_begin,k_end) = rows to produce, [own_lo,own_hi) = owned rows (3, ny-3); no scratch needed. */
int adv_burgers2d_stage_f64(const adv_stage_args* a, const double* u_in, const double* uo, double* u_out, void* cuda_stream);
int adv_burgers2d_stage_f32(const adv_stage_args* a, const float* u_in, const float* uo, float* u_out, void* cuda_stream);

/* No counterpart in the reference.  adv_burgers3d_stage_f64 for ADV_TARGET_BURGERS3D_MGPU without the
 * rowacc argument: the scratch is a fixed array inside the library's device image, so this entry point
 * allocates nothing (no cudaMalloc/cudaFree, no allocating library call) -- for hosts that cannot allocate
 * device memory where they call us: the device budget belongs to the embedding application's pool, the
 * loop cannot take the device-wide synchronisation an allocation implies, or the caller has no allocator.
 * Inputs and outputs are still the caller's buffers.
 *  - Limit: pitch * nzs <= ADV_NOALLOC_MAX_ELEMS (9 MiB of doubles, enough for 1024 x (1024+6) planes);
 *    larger fields are refused with ADV_ERR_INVALID and a message -- use the rowacc variant for them.
 *  - Every GPU that loads the library carries those 9 MiB until the process exits.
 *  - The first call on each device looks up the array's device address.  Make that first call before
 *    starting any stream capture.  The device the stream belongs to must be the CURRENT device on every
 *    call, not only the first (a mismatch is refused with ADV_ERR_INVALID); the looked-up addresses do not
 *    survive cudaDeviceReset.  Host threads may call concurrently for different devices.
 *  - All calls on one device share the one array and it keeps what the last call left: stage 1 clears
 *    the rows of the planes [k_begin,k_end) it produces, stages 2 and 3 continue from stage 1 of the same step.  Two time steps in flight on
 *    the same GPU at once (different streams, or different fields interleaved stage by stage) race on it;
 *    run them one after the other. */
#define ADV_NOALLOC_MAX_ELEMS 1179648
int adv_burgers3d_stage_f64_noalloc(const adv_stage_args* a, const double* u_in, const double* uo, double* u_out, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* ADVDIFF_H */
