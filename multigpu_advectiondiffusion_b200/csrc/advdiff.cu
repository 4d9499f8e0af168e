// libadvdiff.so -- C ABI (include/advdiff.h) over the fused sm_100a stage kernels.
// Host side: plain C-callable functions, no torch, no CPU compute path: every entry point that
// steps a field launches CUDA kernels and fails with ADV_ERR_NO_DEVICE / ADV_ERR_CUDA otherwise.
#include "../../include/advdiff.h"

#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <new>
#include <stdint.h>
#include <thread>
#include <vector>

#include "burgers_kernels.cuh"
#include "burgers2d_kernels.cuh"
#include "burgers_pair_kernels.cuh"
#include "reduce.cuh"
#include "common.cuh"
#include "halo.cuh"
#include "heat_kernels.cuh"

using namespace adv;

// ------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

#define CUDA_TRY(expr)                                                                         \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess)                                                                 \
            return fail(_e == cudaErrorNoDevice || _e == cudaErrorInsufficientDriver           \
                            ? ADV_ERR_NO_DEVICE : ADV_ERR_CUDA,                                \
                        "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

extern "C" const char* adv_last_error(void) { return g_err; }

// A/B switches (ADV_* environment variables) are read at every call, not latched at first use: a caller or a
// test that sets one after the library has already launched something gets what it asked for.
// SM count of the current device (148 on B200; queried, not assumed)
static int current_sm_count()
{
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms > 0 ? sms : 148;
}

static bool env_on(const char* name) { return getenv(name) != nullptr; }
static int env_int(const char* name, int dflt) { const char* v = getenv(name); return v ? atoi(v) : dflt; }
extern "C" int adv_version(void) { return ADV_VERSION; }

extern "C" int adv_device_count(int* count)
{
    if (!count) return fail(ADV_ERR_INVALID, "count is NULL");
    int n = 0, ok = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { *count = 0; return fail(ADV_ERR_NO_DEVICE, "cudaGetDeviceCount: %s", cudaGetErrorString(e)); }
    for (int d = 0; d < n; d++) {
        int major = 0;
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, d) == cudaSuccess && major == 10) ok++;
    }
    *count = ok;
    return ADV_OK;
}

// ------------------------------------------------------------------------------------------
// stage launches
// ------------------------------------------------------------------------------------------
static bool is_heat(int t)
{
    return t == ADV_TARGET_HEAT2D_SGPU || t == ADV_TARGET_HEAT3D_SGPU_BASELINE ||
           t == ADV_TARGET_HEAT3D_SGPU_BLOCKING || t == ADV_TARGET_HEAT3D_MGPU || t == ADV_TARGET_HEAT2D_MGPU;
}
// the two y-slab programs: the slab unit is one ROW (pitch elements), not one z-plane
static bool is_2d_mgpu(int t) { return t == ADV_TARGET_HEAT2D_MGPU || t == ADV_TARGET_BURGERS2D_MGPU; }
static bool is_mgpu(int t) { return t == ADV_TARGET_HEAT3D_MGPU || t == ADV_TARGET_BURGERS3D_MGPU || is_2d_mgpu(t); }
// SURVEY 8f-4 tail: the single-GPU Burgers program (WENO5-Z + viscous term, frozen frame of 3) on z-slabs.  Its array
// has no pad planes, but its frame planes 0..2 and nz-3..nz-1 are never updated, so they play the pads' part: rank r of
// P owns global planes [3 + r*m, 3 + (r+1)*m), m = (nz-6)/P, and stores planes [r*m, r*m + m + 6) -- the layout of the
// MGPU targets with nz-6 "physical" planes.  Parity target: the single-GPU program's own semantics, which the slab
// run reproduces bit for bit for every P (tests/mp_slab_worker.py).
static bool is_sgpu_slab(const adv_desc& d) { return d.target == ADV_TARGET_BURGERS3D_SGPU && d.nranks > 1; }

// second destination of a stage launch (fused halo exchange); internal, not part of the public stage ABI
struct PeerOut {
    void* out2; long long off2;
    const double* cfl_dt; const int* cfl_done; unsigned long long* cfl_umax;
    const SlabSched* fused;     // one launch per stage: boundary windows + flags (interior window = the stage args' [k_begin,k_end))
};
static const PeerOut kNoPeer = {nullptr, 0, nullptr, nullptr, nullptr, nullptr};

// launch schedule of a stage: the caller's fused boundary windows (if any) + the interior window in chunks
static SlabSched make_sched(const SlabSched* fused, int kb, int ke, int chunk)
{
    SlabSched sc;
    if (fused) sc = *fused; else memset(&sc, 0, sizeof(sc));
    sc.ib = kb; sc.ie = ke; sc.chunk = chunk;
    return sc;
}

static int check_stage_args(const adv_stage_args* a, const void* u_in, const void* uo, const void* u_out, int min_n)
{
    if (!a || !u_in || !u_out) return fail(ADV_ERR_INVALID, "NULL argument");
    if (a->stage < 1 || a->stage > 3) return fail(ADV_ERR_INVALID, "stage must be 1, 2 or 3 (got %d)", a->stage);
    if (a->stage > 1 && !uo) return fail(ADV_ERR_INVALID, "uo is required for stages 2 and 3");
    if (u_out == u_in || (a->stage > 1 && u_out == uo)) return fail(ADV_ERR_INVALID, "u_out must not alias an input");
    if (a->nx < min_n || a->ny < min_n) return fail(ADV_ERR_INVALID, "grid %dx%d smaller than the stencil needs (%d)", a->nx, a->ny, min_n);
    if (a->pitch < (size_t)a->nx) return fail(ADV_ERR_INVALID, "pitch %zu < nx %d", a->pitch, a->nx);
    return ADV_OK;
}

template <typename REAL>
static int heat_args_from(const adv_stage_args* a, HeatArgs<REAL>& h)
{
    h.out2 = nullptr; h.off2 = 0;
    h.nx = a->nx; h.ny = a->ny; h.nzs = a->nzs;
    h.pitch = (long long)a->pitch; h.plane = (long long)a->pitch * a->ny;
    h.kb = a->k_begin; h.ke = a->k_end;
    h.stage = a->stage;
    h.kx = (REAL)a->kx; h.ky = (REAL)a->ky; h.kz = (REAL)a->kz; h.dt = (REAL)a->dt;
    h.iso = 0;
    const int nx = a->nx, ny = a->ny, nz = a->nzs;
    switch (a->target) {
    case ADV_TARGET_HEAT2D_SGPU:            // kernels.cu:53,83
        h.lap = Box3{{3, 3, 0}, {nx - 4, ny - 4, 0}};
        h.rk = h.lap;
        break;
    case ADV_TARGET_HEAT3D_SGPU_BASELINE:   // kernels.cu:94,127
        h.lap = Box3{{2, 2, 2}, {nx - 3, ny - 3, nz - 3}};
        h.rk = h.lap;
        break;
    case ADV_TARGET_HEAT3D_SGPU_BLOCKING:   // kernels.cu:59-85 (3, nz-2, 16) ; :113
        h.lap = Box3{{3, 3, 3}, {nx - 4, ny - 4, nz - 3}};
        h.rk = Box3{{2, 2, 2}, {nx - 3, ny - 3, nz - 3}};
        break;
    case ADV_TARGET_HEAT3D_MGPU:            // Kernels.cu:236,288
        h.lap = Box3{{3, 3, a->own_lo}, {nx - 4, ny - 4, a->own_hi - 1}};
        h.rk = Box3{{3, 3, 0}, {nx - 4, ny - 4, nz - 1}};
        h.iso = 1;
        break;
    case ADV_TARGET_HEAT2D_MGPU:            // MultiGPU/Diffusion2d_Baseline/Kernels.cu:144-151,170 (ny = stored rows)
        // Compute_Laplace starts one row below its window (j = ... + jstart-1): on the first slab that is pad row 2
        h.lap = Box3{{3, a->own_lo - (a->k_global0 == 0 ? 1 : 0), 0}, {nx - 4, a->own_hi - 1, 0}};
        h.rk = Box3{{0, 0, 0}, {nx - 1, ny - 1, 0}};
        break;
    default:
        return fail(ADV_ERR_INVALID, "target %d is not a heat target", a->target);
    }
    return ADV_OK;
}

// cuTensorMapEncodeTiled through the runtime's driver-entry-point lookup (no libcuda link dependency)
typedef CUresult (*tmap_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static tmap_encode_fn lookup_tmap_encode()
{
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
        return (tmap_encode_fn)p;
    return nullptr;
}
static tmap_encode_fn get_tmap_encode()
{
    // function-local static: initialised exactly once even when the ranks are threads of one process (a
    // hand-rolled "tried" flag let a second thread see "not available" and pick a different halo path)
    static const tmap_encode_fn fn = lookup_tmap_encode();
    return fn;
}

static int encode_tmap3(CUtensorMap* tm, int esz, const void* base, int nx, int ny, int nz, size_t pitch, int bx, int by)
{
    tmap_encode_fn enc = get_tmap_encode();
    if (!enc) return fail(ADV_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
    const cuuint64_t gdim[3] = {(cuuint64_t)nx, (cuuint64_t)ny, (cuuint64_t)nz};
    const cuuint64_t gstr[2] = {(cuuint64_t)pitch * esz, (cuuint64_t)pitch * ny * esz};
    const cuuint32_t box[3] = {(cuuint32_t)bx, (cuuint32_t)by, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(tm, esz == 8 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)base,
                     gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(ADV_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    return ADV_OK;
}

template <typename REAL, int TY, int NST>
static int heat3d_launch_tma(const adv_stage_args* a, const HeatArgs<REAL>& h, const REAL* u_in, const REAL* uo, REAL* u_out, cudaStream_t st,
                             const SlabSched* fused = nullptr)
{
    typedef HeatTmaSmem<REAL, TY, NST> S;
    CUtensorMap tm, tmo;
    int rc = encode_tmap3(&tm, sizeof(REAL), u_in, a->nx, a->ny, a->nzs, a->pitch, S::BX, S::BY);
    if (rc) return rc;
    rc = encode_tmap3(&tmo, sizeof(REAL), a->stage == 1 ? u_in : uo, a->nx, a->ny, a->nzs, a->pitch, S::TX, TY);
    if (rc) return rc;
    const int planes = a->k_end - a->k_begin;
    const int gx = (a->nx + S::TX - 1) / S::TX, gy = (a->ny + TY - 1) / TY;
    // z chunks: equal-length CTAs run in waves of (resident CTAs per SM) x (SMs); a launch costs about
    // waves x (chunk + 4 run-in planes).  Pick the chunk count that minimises that (ADV_HEAT_NCH forces one;
    // ADV_HEAT_CHUNK_OLD=1 is the round-1 rule: ~2368 CTAs, chunks of >= 16 planes -- 4.6 waves at 256^3).
    const int nch_env = env_int("ADV_HEAT_NCH", 0);
    int nch = (2368 + gx * gy - 1) / (gx * gy);
    if (nch > planes / 16) nch = planes / 16;
    if (nch < 1) nch = 1;
    if (!env_on("ADV_HEAT_CHUNK_OLD") && planes > 0) {
        int sms = 148, per_sm = 3, dev = 0;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        {
            auto k2 = heat3d_stage_tma<REAL, TY, NST, 2>;
            if (sizeof(S) > 48 * 1024) cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(S));
            int occ = 0;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k2, 32 * (TY + 1), sizeof(S)) == cudaSuccess && occ > 0) per_sm = occ;
        }
        const long long slots = (long long)sms * per_sm;
        long long best = -1;
        // chunks of 16..48 planes (measured at 512^3: 22..44 planes give 96.3-97.0 % of the copy peak, 87 planes
        // 94.3 %, 14 planes 94.6 %; the 4 run-in planes of a chunk are L2 hits and cost about 2 planes' time);
        // among those, the count with the fewest waves x (chunk + 2): at 256^3 that is 10 chunks (2.9 waves, 85 %)
        // where 9 or 11 chunks give 82 % and 80 %
        for (int c = 1; c <= planes / 16 || c == 1; c++) {
            const int ch = (planes + c - 1) / c, nz = (planes + ch - 1) / ch;
            if (ch > 48 && planes / 16 >= 2 && (planes + c) / (c + 1) >= 16) continue;
            const long long ctas = (long long)gx * gy * nz, waves = (ctas + slots - 1) / slots;
            const long long cost = waves * (ch + 2);
            if (best < 0 || cost < best) { best = cost; nch = c; }
        }
    }
    if (nch_env > 0) nch = nch_env;
    const int chunk = planes > 0 ? (planes + nch - 1) / nch : 1;
    const SlabSched sc = make_sched(fused, a->k_begin, a->k_end, chunk);
    dim3 grid(gx, gy, sc.nb + (planes + chunk - 1) / chunk), block(32, TY + 1, 1);
    const size_t shm = sizeof(S);
#define HEAT_TMA_LAUNCH(STAGE_) do { \
        auto kfn = heat3d_stage_tma<REAL, TY, NST, STAGE_>; \
        if (shm > 48 * 1024) CUDA_TRY(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm)); \
        kfn<<<grid, block, shm, st>>>(tm, tmo, h, sc, u_out); } while (0)
    if (a->stage == 1) HEAT_TMA_LAUNCH(1); else if (a->stage == 2) HEAT_TMA_LAUNCH(2); else HEAT_TMA_LAUNCH(3);
#undef HEAT_TMA_LAUNCH
    CUDA_TRY(cudaGetLastError());
    return ADV_OK;
}

template <typename REAL>
static bool heat3d_tma_ok(const adv_stage_args* a, const REAL* u_in, const REAL* uo, const REAL* u_out)
{
    const size_t vec = 16 / sizeof(REAL);
    return (a->nx % 2 == 0) && (a->pitch % vec == 0) && a->nx >= 8 && ((uintptr_t)u_in % 16 == 0) && ((uintptr_t)u_out % 16 == 0) &&
           (a->stage == 1 || (uintptr_t)uo % 16 == 0) && get_tmap_encode() != nullptr;
}

template <typename REAL>
static int heat3d_stage_impl(const adv_stage_args* a, const REAL* u_in, const REAL* uo, REAL* u_out, cudaStream_t st, PeerOut peer = kNoPeer)
{
    int rc = check_stage_args(a, u_in, uo, u_out, 5);
    if (rc) return rc;
    if (a->k_begin < 0 || a->k_end > a->nzs || a->k_begin > a->k_end) return fail(ADV_ERR_INVALID, "bad plane window [%d,%d) of %d", a->k_begin, a->k_end, a->nzs);
    if (a->k_begin == a->k_end && !peer.fused) return ADV_OK;
    HeatArgs<REAL> h;
    rc = heat_args_from<REAL>(a, h);
    if (rc) return rc;
    h.out2 = (REAL*)peer.out2; h.off2 = peer.off2;
    // TMA path: needs 16-byte rows and pairs of cells; anything else takes the direct kernel
    const int heat_variant = env_on("ADV_HEAT_DIRECT") ? 0 : 1;     // A/B switch: ADV_HEAT_DIRECT=1 forces the direct kernel
    const bool tma_ok = heat_variant > 0 && heat3d_tma_ok<REAL>(a, u_in, uo, u_out);
    if ((peer.out2 || peer.fused) && !tma_ok) return fail(ADV_ERR_INVALID, "fused halo stores need the TMA heat kernel (even nx, 16-byte rows)");
    if (tma_ok) return heat3d_launch_tma<REAL, 8, 6>(a, h, u_in, uo, u_out, st, peer.fused);     // 64x8 tile, 6 planes in flight
    constexpr int CHUNK = 32;
    dim3 block(32, 8, 1);
    dim3 grid((a->nx + 31) / 32, (a->ny + 7) / 8, (a->k_end - a->k_begin + CHUNK - 1) / CHUNK);
    heat3d_stage_direct<REAL, CHUNK><<<grid, block, 0, st>>>(h, u_in, a->stage == 1 ? u_in : uo, u_out);
    CUDA_TRY(cudaGetLastError());
    return ADV_OK;
}

// EXPERIMENTAL (ADV_HEAT2D_TMA=1, off by default; written after this round's GPU time was spent and not yet run
// on hardware): the 3-D TMA pipeline applied to a 2-D field seen as nx x 1 x rows -- one consumer warp and one
// producer lane per 64-column strip, 8 rows in flight per strip.  Same arithmetic as heat2d_stage_rows.
template <typename REAL>
static int heat2d_launch_tma(const adv_stage_args* a, const HeatArgs<REAL>& h2d, const REAL* u_in, const REAL* uo, REAL* u_out, cudaStream_t st)
{
    constexpr int TY = 1, NST = 8;
    typedef HeatTmaSmem<REAL, TY, NST, true> S;
    const int rows_total = a->ny;
    HeatArgs<REAL> h = h2d;
    h.ny = 1; h.nzs = rows_total; h.plane = h.pitch;
    h.lap = Box3{{h2d.lap.lo[0], 0, h2d.lap.lo[1]}, {h2d.lap.hi[0], 0, h2d.lap.hi[1]}};
    h.rk = Box3{{h2d.rk.lo[0], 0, h2d.rk.lo[1]}, {h2d.rk.hi[0], 0, h2d.rk.hi[1]}};
    h.kz = h2d.ky; h.ky = REAL(0); h.iso = 0;
    CUtensorMap tm, tmo;
    int rc = encode_tmap3(&tm, sizeof(REAL), u_in, a->nx, 1, rows_total, a->pitch, S::BX, S::BY);
    if (rc) return rc;
    rc = encode_tmap3(&tmo, sizeof(REAL), a->stage == 1 ? u_in : uo, a->nx, 1, rows_total, a->pitch, S::TX, TY);
    if (rc) return rc;
    const int rows = h.ke - h.kb, gx = (a->nx + S::TX - 1) / S::TX;
    int nch = (8 * 1184 + gx - 1) / gx;                 // ~8 waves of 8 resident strips per SM
    if (nch > rows / 32) nch = rows / 32;               // chunks of >= 32 rows (4 run-in rows each)
    if (nch < 1) nch = 1;
    const int chunk = (rows + nch - 1) / nch;
    dim3 grid(gx, 1, (rows + chunk - 1) / chunk), block(32, TY + 1, 1);
    const size_t shm = sizeof(S);
#define HEAT2D_TMA_LAUNCH(STAGE_) do { \
        auto kfn = heat3d_stage_tma<REAL, TY, NST, STAGE_, true>; \
        kfn<<<grid, block, shm, st>>>(tm, tmo, h, make_sched(nullptr, h.kb, h.ke, chunk), u_out); } while (0)
    if (a->stage == 1) HEAT2D_TMA_LAUNCH(1); else if (a->stage == 2) HEAT2D_TMA_LAUNCH(2); else HEAT2D_TMA_LAUNCH(3);
#undef HEAT2D_TMA_LAUNCH
    CUDA_TRY(cudaGetLastError());
    return ADV_OK;
}

template <typename REAL>
static int heat2d_stage_impl(const adv_stage_args* a, const REAL* u_in, const REAL* uo, REAL* u_out, cudaStream_t st, PeerOut peer = kNoPeer)
{
    int rc = check_stage_args(a, u_in, uo, u_out, 7);
    if (rc) return rc;
    const bool slab = a->target == ADV_TARGET_HEAT2D_MGPU;
    if (slab && (a->k_begin < 0 || a->k_end > a->ny || a->k_begin > a->k_end)) return fail(ADV_ERR_INVALID, "bad row window [%d,%d) of %d", a->k_begin, a->k_end, a->ny);
    if (slab && (a->own_lo < 3 || a->own_hi > a->ny - 3)) return fail(ADV_ERR_INVALID, "owned rows [%d,%d) need 3 rows of halo inside [0,%d)", a->own_lo, a->own_hi, a->ny);
    HeatArgs<REAL> h;
    adv_stage_args b = *a;
    b.nzs = 1; b.k_begin = 0; b.k_end = 1;
    rc = heat_args_from<REAL>(&b, h);
    if (rc) return rc;
    h.kb = slab ? a->k_begin : 0; h.ke = slab ? a->k_end : a->ny;       // ROW window for the 2-D kernels
    if (h.kb == h.ke) return ADV_OK;
    h.out2 = (REAL*)peer.out2; h.off2 = peer.off2;
    const size_t vb = 2 * sizeof(REAL);
    const bool vec_ok = a->pitch % 2 == 0 && (uintptr_t)u_in % vb == 0 && (uintptr_t)u_out % vb == 0 && (a->stage == 1 || (uintptr_t)uo % vb == 0) &&
                        (!peer.out2 || ((uintptr_t)peer.out2 % vb == 0 && peer.off2 % 2 == 0));
    const bool heat2d_tma = env_on("ADV_HEAT2D_TMA");     // experimental, see heat2d_launch_tma
    if (heat2d_tma && vec_ok && a->nx % 2 == 0 && a->nx >= 8 && a->pitch % (16 / sizeof(REAL)) == 0 &&
        (uintptr_t)u_in % 16 == 0 && (uintptr_t)u_out % 16 == 0 && (a->stage == 1 || (uintptr_t)uo % 16 == 0) && get_tmap_encode())
        return heat2d_launch_tma<REAL>(a, h, u_in, uo, u_out, st);
    // pairs-of-cells row-marching kernel for slabs and large grids; one-thread-per-cell kernel for the small
    // single-GPU case (512^2 lives in L2 and is launch-bound) and for unaligned caller buffers
    if (vec_ok && (slab || (size_t)a->nx * a->ny > ((size_t)1 << 21))) {
        const int rows = h.ke - h.kb, gx = (a->nx / 2 + 1 + 127) / 128;
        const int ctas_per_sm = env_int("ADV_HEAT2D_CTAS", 128);  // A/B switch; measured at 8192^2: 16 -> 184, 64 -> 239, 128 -> 253 Gcell-updates/s (equal-length CTAs: many waves keep the tail short)
        int nch = (ctas_per_sm * current_sm_count() + gx - 1) / gx;        // many short chunks: loads in flight matter more than the 4 run-in rows ...
        if (nch > rows / 16) nch = rows / 16;               // ... but chunks of >= 16 rows (the 4 run-in rows are L2 hits: neighbouring chunks run together)
        if (nch < 1) nch = 1;
        const int chunk = (rows + nch - 1) / nch;
        dim3 grid(gx, (rows + chunk - 1) / chunk, 1), block(128, 1, 1);
        const REAL* uo_ = a->stage == 1 ? u_in : uo;
#define LAUNCH_ROWS(STAGE_) do { \
            if (peer.out2) heat2d_stage_rows<REAL, STAGE_, true><<<grid, block, 0, st>>>(h, chunk, u_in, uo_, u_out); \
            else heat2d_stage_rows<REAL, STAGE_, false><<<grid, block, 0, st>>>(h, chunk, u_in, uo_, u_out); } while (0)
        if (a->stage == 1) LAUNCH_ROWS(1); else if (a->stage == 2) LAUNCH_ROWS(2); else LAUNCH_ROWS(3);
#undef LAUNCH_ROWS
    } else {
        dim3 block(32, 8, 1);
        dim3 grid((a->nx + 31) / 32, (h.ke - h.kb + 7) / 8, 1);
        heat2d_stage_direct<REAL><<<grid, block, 0, st>>>(h, u_in, a->stage == 1 ? u_in : uo, u_out);
    }
    CUDA_TRY(cudaGetLastError());
    return ADV_OK;
}

// Burgers 2-D y-slab stage (MultiGPU/Burgers2d_Baseline): ny = stored rows, [k_begin,k_end) = row window
template <typename REAL>
static int burgers2d_stage_impl(const adv_stage_args* a, const REAL* u_in, const REAL* uo, REAL* u_out, cudaStream_t st, PeerOut peer = kNoPeer)
{
    int rc = check_stage_args(a, u_in, uo, u_out, 8);
    if (rc) return rc;
    if (a->target != ADV_TARGET_BURGERS2D_MGPU) return fail(ADV_ERR_INVALID, "target %d is not the 2-D Burgers target", a->target);
    if (a->k_begin < 0 || a->k_end > a->ny || a->k_begin > a->k_end) return fail(ADV_ERR_INVALID, "bad row window [%d,%d) of %d", a->k_begin, a->k_end, a->ny);
    if (a->own_lo < 3 || a->own_hi > a->ny - 3) return fail(ADV_ERR_INVALID, "owned rows [%d,%d) need 3 rows of halo inside [0,%d)", a->own_lo, a->own_hi, a->ny);
    if (a->k_begin == a->k_end) return ADV_OK;
    BurgersArgs<REAL> b;
    memset(&b, 0, sizeof(b));
    b.nx = a->nx; b.ny = a->ny; b.nzs = 1;
    b.pitch = (long long)a->pitch; b.plane = (long long)a->pitch * a->ny;
    b.kb = a->k_begin; b.ke = a->k_end; b.own_lo = a->own_lo; b.own_hi = a->own_hi;
    b.stage = a->stage;
    b.dt = (REAL)a->dt; b.dx = (REAL)a->dx; b.dy = (REAL)a->dy;
    b.rdx = REAL(1) / b.dx; b.rdy = REAL(1) / b.dy;
    b.rdx6 = b.rdx / REAL(6); b.rdy6 = b.rdy / REAL(6);
    b.out2 = (REAL*)peer.out2; b.off2 = peer.off2;
    b.quirk_lo = 1; b.quirk_hi = 0;
    fill_weno_consts<REAL>(b, WENO_JS);
    constexpr int TX = 64, TY = 16, NT = 256;
    dim3 grid((a->nx + TX - 1) / TX, (a->k_end - a->k_begin + TY - 1) / TY, 1), block(NT, 1, 1);
    const REAL* uo_ = a->stage == 1 ? u_in : uo;
    if (env_on("ADV_BURGERS_DIRECT")) {     // parity anchor: reference operand order, one thread per cell
        dim3 dblock(64, 4, 1), dgrid((a->nx + 63) / 64, (a->k_end - a->k_begin + 3) / 4, 1);
        if (peer.out2) burgers2d_stage_direct<REAL, WENO_JS, true><<<dgrid, dblock, 0, st>>>(b, u_in, uo_, u_out);
        else burgers2d_stage_direct<REAL, WENO_JS, false><<<dgrid, dblock, 0, st>>>(b, u_in, uo_, u_out);
        CUDA_TRY(cudaGetLastError());
        return ADV_OK;
    }
    const size_t shm = sizeof(Burgers2dSmem<REAL>);
#define LAUNCH_B2D(STAGE_, PEER_) do { \
        auto kfn = burgers2d_stage_tile<REAL, WENO_JS, STAGE_, PEER_>; \
        CUDA_TRY(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm)); \
        kfn<<<grid, block, shm, st>>>(b, u_in, uo_, u_out); } while (0)
#define LAUNCH_B2D_S(PEER_) do { \
        if (a->stage == 1) LAUNCH_B2D(1, PEER_); else if (a->stage == 2) LAUNCH_B2D(2, PEER_); else LAUNCH_B2D(3, PEER_); } while (0)
    if (peer.out2) LAUNCH_B2D_S(true); else LAUNCH_B2D_S(false);
#undef LAUNCH_B2D_S
#undef LAUNCH_B2D
    CUDA_TRY(cudaGetLastError());
    return ADV_OK;
}

template <typename REAL>
static int burgers3d_stage_impl(const adv_stage_args* a, const REAL* u_in, const REAL* uo, REAL* u_out, REAL* rowacc, cudaStream_t st, PeerOut peer = kNoPeer)
{
    int rc = check_stage_args(a, u_in, uo, u_out, 7);
    if (rc) return rc;
    if (a->k_begin < 0 || a->k_end > a->nzs || a->k_begin > a->k_end) return fail(ADV_ERR_INVALID, "bad plane window [%d,%d) of %d", a->k_begin, a->k_end, a->nzs);
    if (a->own_lo < 3 || a->own_hi > a->nzs - 3) return fail(ADV_ERR_INVALID, "owned planes [%d,%d) need 3 planes of halo inside [0,%d)", a->own_lo, a->own_hi, a->nzs);
    if (a->k_begin == a->k_end && !peer.fused) return ADV_OK;
    const bool mg = a->target == ADV_TARGET_BURGERS3D_MGPU;
    if (!mg && a->target != ADV_TARGET_BURGERS3D_SGPU) return fail(ADV_ERR_INVALID, "target %d is not a Burgers target", a->target);
    if (mg && !rowacc) return fail(ADV_ERR_INVALID, "rowacc scratch is required for ADV_TARGET_BURGERS3D_MGPU");

    BurgersArgs<REAL> b;
    memset(&b, 0, sizeof(b));
    b.cfl_dt = peer.cfl_dt; b.cfl_done = peer.cfl_done; b.cfl_umax = peer.cfl_umax;
    b.nx = a->nx; b.ny = a->ny; b.nzs = a->nzs;
    b.pitch = (long long)a->pitch; b.plane = (long long)a->pitch * a->ny;
    b.kb = a->k_begin; b.ke = a->k_end; b.own_lo = a->own_lo; b.own_hi = a->own_hi;
    b.stage = a->stage;
    b.dt = (REAL)a->dt; b.dx = (REAL)a->dx; b.dy = (REAL)a->dy; b.dz = (REAL)a->dz;
    b.rdx = REAL(1) / b.dx; b.rdy = REAL(1) / b.dy; b.rdz = REAL(1) / b.dz;
    b.rdx6 = b.rdx / REAL(6); b.rdy6 = b.rdy / REAL(6); b.rdz6 = b.rdz / REAL(6);
    b.kx = (REAL)a->kx; b.ky = (REAL)a->ky; b.kz = (REAL)a->kz;
    b.out2 = (REAL*)peer.out2; b.off2 = peer.off2;
    b.visc = (!mg && (a->kx != 0 || a->ky != 0 || a->kz != 0)) ? 1 : 0;
    b.quirk_lo = 1; b.quirk_hi = 0;
    if (!mg && !(a->flags & ADV_FLAG_VISC_ONCE)) { b.quirk_lo = 16 - a->k_global0; b.quirk_hi = 18 - a->k_global0; }

    constexpr int CHUNK = 16;
    dim3 block(32, 4, 1);
    dim3 grid((a->nx + 31) / 32, (a->ny + 3) / 4, (a->k_end - a->k_begin + CHUNK - 1) / CHUNK);
    const bool force_direct = env_on("ADV_BURGERS_DIRECT");   // A/B switch for tests and profiling
    fill_weno_consts<REAL>(b, mg ? WENO_JS : WENO_Z);
    // default: the pair kernel (TMA-staged planes; needs 16-byte rows for the tensor map)
    const bool pair_ok = !force_direct && !env_on("ADV_BURGERS_RING") && get_tmap_encode() != nullptr &&
                         ((uintptr_t)u_in % 16 == 0) && ((a->pitch * sizeof(REAL)) % 16 == 0) && a->nx >= 8;
    if (!mg && peer.out2 && !pair_ok) return fail(ADV_ERR_STATE, "peer-store halos of the single-GPU Burgers semantics need the pair kernel");
    if ((peer.cfl_done || peer.fused) && !pair_ok) return fail(ADV_ERR_STATE, "the device-resident time step and the one-launch stage need the pair kernel");
    if (pair_ok) {
        typedef BurgersPairSmem<REAL> S;
        CUtensorMap tm;
        rc = encode_tmap3(&tm, sizeof(REAL), u_in, a->nx, a->ny, a->nzs, a->pitch, S::RP, S::ROWS);
        if (rc) return rc;
        const int planes = a->k_end - a->k_begin;
        const int gx = (a->nx + S::TX - 1) / S::TX, gy = (a->ny + S::TY - 1) / S::TY;
        int sms = 148;
        { int dev = 0; if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
        // z chunks: all CTAs take the same time, so the launch runs in waves of 2 CTAs per SM and costs about
        // waves x (chunk + 3) plane-times (5 planes of staging and one plane of z halves are the run-in).  Measured at
        // 512^3 (tools/r02_ab_nch.sh): 10-12 chunks (44-52 planes) 7.26 ms per step, 4 chunks (130 planes, the
        // fewest waves) 7.37, 16 chunks 7.35, 37 chunks (14 planes) 8.02 -- many mid-sized CTAs balance better than the
        // wave count alone says.  So: chunks of 16..64 planes where the slab allows, and among those the fewest
        // waves x (chunk + 4): 9 chunks of 58 planes at 512^3 (7.19 ms), 9 chunks of 30 planes at 256^3 (four full waves;
        // a 32-plane floor picked 6 chunks = 2.6 waves there and lost 6 %).
        const int minchunk = env_int("ADV_BURGERS_MINCHUNK", 12), forced = env_int("ADV_BURGERS_NCH", 0);
        int nch = 1;
        long long best = -1;
        for (int c = 1; c <= planes / minchunk || c == 1; c++) {
            const int ch = (planes + c - 1) / c, nz = (planes + ch - 1) / ch;
            if (planes >= 64 && (ch > 64 || ch < 16)) continue;
            const long long ctas = (long long)gx * gy * nz, waves = (ctas + 2 * sms - 1) / (2 * sms);
            const long long cost = waves * (ch + 4);
            if (best < 0 || cost < best) { best = cost; nch = c; }
        }
        if (forced > 0) nch = forced;
        const int chunk = planes > 0 ? (planes + nch - 1) / nch : 1;
        const SlabSched sc = make_sched(peer.fused, a->k_begin, a->k_end, chunk);
        dim3 pgrid(gx, gy, sc.nb + (planes + chunk - 1) / chunk), pblock(32, S::NW, 1);
        const REAL* uo_ = a->stage == 1 ? u_in : uo;
        const size_t shm = sizeof(S);
#define LAUNCH_PAIR(W_, M_, STAGE_, PEER_, TRACK_) do { \
            auto kfn = burgers3d_stage_pair<REAL, W_, M_, STAGE_, PEER_, TRACK_>; \
            CUDA_TRY(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm)); \
            kfn<<<pgrid, pblock, shm, st>>>(tm, b, sc, u_in, uo_, u_out, rowacc); } while (0)
#define LAUNCH_PAIR_S(W_, M_, PEER_) do { \
            if (a->stage == 1) LAUNCH_PAIR(W_, M_, 1, PEER_, false); \
            else if (a->stage == 2) LAUNCH_PAIR(W_, M_, 2, PEER_, false); \
            else if (b.cfl_umax) LAUNCH_PAIR(W_, M_, 3, PEER_, true);   /* stage 3 of a dynamic CFL run also tracks max|u| */ \
            else LAUNCH_PAIR(W_, M_, 3, PEER_, false); } while (0)
        if (!mg && (b.out2 || peer.fused)) LAUNCH_PAIR_S(WENO_Z, BMODE_SGPU, true);
        else if (!mg) LAUNCH_PAIR_S(WENO_Z, BMODE_SGPU, false);
        else if (b.out2 || peer.fused) LAUNCH_PAIR_S(WENO_JS, BMODE_MGPU, true);
        else LAUNCH_PAIR_S(WENO_JS, BMODE_MGPU, false);
#undef LAUNCH_PAIR_S
#undef LAUNCH_PAIR
    } else if (!force_direct) {
        // ring kernel (ADV_BURGERS_RING=1, or rows that are not 16-byte aligned): 32x16 tile, 2 cells per thread, 8+1 warps, 2 CTAs/SM
        constexpr int TY = 16, CPT = 2, MINB = 2;
        const int planes = a->k_end - a->k_begin;
        const int gx = (a->nx + 31) / 32, gy = (a->ny + TY - 1) / TY;
        int nch = (12 * current_sm_count() + gx * gy - 1) / (gx * gy);         // ~6 waves of 2 CTAs/SM ...
        if (nch > planes / 16) nch = planes / 16;           // ... but chunks of >= 16 planes (6 planes of z run-in each)
        if (nch < 1) nch = 1;
        const int chunk = (planes + nch - 1) / nch;
        dim3 tgrid(gx, gy, (planes + chunk - 1) / chunk), tblock(32, TY / CPT + 1, 1);
        const REAL* uo_ = a->stage == 1 ? u_in : uo;
        const size_t shm = sizeof(BurgersRingSmem<REAL, TY>);
#define LAUNCH_RING(W_, M_, STAGE_, PEER_) do { \
            auto kfn = burgers3d_stage_ring<REAL, W_, M_, TY, CPT, MINB, STAGE_, PEER_>; \
            CUDA_TRY(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm)); \
            kfn<<<tgrid, tblock, shm, st>>>(b, chunk, u_in, uo_, u_out, rowacc); } while (0)
#define LAUNCH_RING_S(W_, M_, PEER_) do { \
            if (a->stage == 1) LAUNCH_RING(W_, M_, 1, PEER_); \
            else if (a->stage == 2) LAUNCH_RING(W_, M_, 2, PEER_); \
            else LAUNCH_RING(W_, M_, 3, PEER_); } while (0)
        if (!mg) LAUNCH_RING_S(WENO_Z, BMODE_SGPU, false);
        else if (b.out2) LAUNCH_RING_S(WENO_JS, BMODE_MGPU, true);
        else LAUNCH_RING_S(WENO_JS, BMODE_MGPU, false);
#undef LAUNCH_RING_S
#undef LAUNCH_RING
    } else if (mg)
        burgers3d_stage_direct<REAL, WENO_JS, BMODE_MGPU, CHUNK><<<grid, block, 0, st>>>(b, u_in, a->stage == 1 ? u_in : uo, u_out, rowacc);
    else
        burgers3d_stage_direct<REAL, WENO_Z, BMODE_SGPU, CHUNK><<<grid, block, 0, st>>>(b, u_in, a->stage == 1 ? u_in : uo, u_out, rowacc);
    CUDA_TRY(cudaGetLastError());
    return ADV_OK;
}

extern "C" int adv_heat3d_stage_f64(const adv_stage_args* a, const double* u_in, const double* uo, double* u_out, void* st)
{ return heat3d_stage_impl<double>(a, u_in, uo, u_out, (cudaStream_t)st); }
extern "C" int adv_heat3d_stage_f32(const adv_stage_args* a, const float* u_in, const float* uo, float* u_out, void* st)
{ return heat3d_stage_impl<float>(a, u_in, uo, u_out, (cudaStream_t)st); }
extern "C" int adv_heat2d_stage_f64(const adv_stage_args* a, const double* u_in, const double* uo, double* u_out, void* st)
{ return heat2d_stage_impl<double>(a, u_in, uo, u_out, (cudaStream_t)st); }
extern "C" int adv_heat2d_stage_f32(const adv_stage_args* a, const float* u_in, const float* uo, float* u_out, void* st)
{ return heat2d_stage_impl<float>(a, u_in, uo, u_out, (cudaStream_t)st); }
extern "C" int adv_burgers2d_stage_f64(const adv_stage_args* a, const double* u_in, const double* uo, double* u_out, void* st)
{ return burgers2d_stage_impl<double>(a, u_in, uo, u_out, (cudaStream_t)st); }
extern "C" int adv_burgers2d_stage_f32(const adv_stage_args* a, const float* u_in, const float* uo, float* u_out, void* st)
{ return burgers2d_stage_impl<float>(a, u_in, uo, u_out, (cudaStream_t)st); }
extern "C" int adv_burgers3d_stage_f64(const adv_stage_args* a, const double* u_in, const double* uo, double* u_out, double* rowacc, void* st)
{ return burgers3d_stage_impl<double>(a, u_in, uo, u_out, rowacc, (cudaStream_t)st); }
extern "C" int adv_burgers3d_stage_f32(const adv_stage_args* a, const float* u_in, const float* uo, float* u_out, float* rowacc, void* st)
{ return burgers3d_stage_impl<float>(a, u_in, uo, u_out, rowacc, (cudaStream_t)st); }

// ------------------------------------------------------------------------------------------
// solver handle
// ------------------------------------------------------------------------------------------
// dynamic-CFL state on the device (second half of the arena's flag block) and what the host sees of it
struct CflDevState {
    double dt, t, it, tEnd, cfl, dx;        // REAL-valued, widened
    int done, max_steps;
    unsigned long long local_max;           // bits of max|u| over what this rank's stage-3 kernels stored
    unsigned long long global_max;          // after the all-reduce over ranks
    unsigned seq;                           // dt kernels executed
};
struct CflHostStatus { volatile unsigned seq; volatile int done; volatile double t, dt, it; };

struct adv_solver {
    adv_desc d;
    size_t esz;                 // sizeof(REAL)
    int nzl;                    // physical slab units (z-planes; rows for the 2-D y-slab targets) owned by this rank
    int nzs;                    // stored units (nzl + 6 for MGPU targets)
    int urows;                  // rows per slab unit: ny for z-slabs, 1 for y-slabs
    int nsplit;                 // physical extent along the split axis (nz, or ny for y-slabs)
    size_t pitch;               // elements
    size_t buf_bytes;
    void* buf[3];               // ping-pong field buffers; buf[cur] holds the solution
    int cur;
    void* rowacc;               // Burgers MGPU scratch (pitch * nzs)
    // one device allocation ("arena") holds the three buffers, the scratch and the halo flags, so that a
    // neighbour rank can map all of it with a single IPC handle and address it by the same offsets
    char* arena;
    size_t arena_bytes, off_buf[3], off_rowacc, off_flags;
    char* peer_base[2];         // [0] left, [1] right neighbour's arena as seen from this device
    bool peer_is_ipc[2];
    int p2p;                    // 1: halo planes are stored straight into the neighbours' ghost planes
    unsigned stage_counter;     // stages executed so far (the value the halo flags carry)
    unsigned probe_counter;     // exchanges run by adv_halo_probe (carried by the probe flags)
    // dynamic CFL run (adv_run_until_cfl): the time-step state lives in the flag block of the arena; the host
    // follows it through a few words of mapped pinned memory that the dt kernel writes
    bool cfl_on;
    struct CflHostStatus* cfl_host;     // mapped pinned
    struct CflHostStatus* cfl_host_dev; // the same words as the device sees them
    // small single-rank problems are launch-bound: three time steps (one full rotation of the three
    // buffers, nine stage launches) are captured once into a CUDA graph and replayed
    cudaGraphExec_t graph_exec;
    int graph_cur;              // value of `cur` the captured graph starts from
    cudaStream_t stream, comm_stream;
    cudaEvent_t ev_start, ev_stop, ev_boundary, ev_comm;
    bool uploaded;
    long long launches;
    // constants in double (computed in REAL then widened, see make_constants)
    double dx, dy, dz, dt_fixed, kx, ky, kz, visc_c;
    HaloComm comm;
};

// The drivers compute their constants in REAL; keep that arithmetic (a5 in SURVEY.md 8a).
template <typename REAL>
static void make_constants(adv_solver* s)
{
    const adv_desc& d = s->d;
    const REAL L = (REAL)d.lx, W = (REAL)d.ly, H = (REAL)d.lz;
    const unsigned Nx = d.nx, Ny = d.ny, Nz = d.nz;
    REAL dx = L / (Nx - 1), dy = W / (Ny - 1), dz = 0;
    REAL dt = 0, kx = 0, ky = 0, kz = 0, C = 0;
    switch (d.target) {
    case ADV_TARGET_HEAT2D_SGPU: {          // SingleGPU/Diffusion2d/main.cpp:53-58
        C = (REAL)d.coef;
        dt = 1 / (2 * C * (1 / dx / dx + 1 / dy / dy)) * 0.9;
        kx = C / (12 * dx * dx); ky = C / (12 * dy * dy);
        break;
    }
    case ADV_TARGET_HEAT3D_SGPU_BASELINE:
    case ADV_TARGET_HEAT3D_SGPU_BLOCKING: { // SingleGPU/Diffusion3d_baselineCode/main.cpp:55-61
        C = (REAL)d.coef;
        dz = H / (Nz - 1);
        dt = 1 / (2 * C * (1 / dx / dx + 1 / dy / dy + 1 / dz / dz)) * 0.9;
        kx = C / (12 * dx * dx); ky = C / (12 * dy * dy); kz = C / (12 * dz * dz);
        break;
    }
    case ADV_TARGET_HEAT3D_MGPU: {          // MultiGPU/Diffusion3d_Baseline/main.c:61-67
        C = (REAL)d.coef;
        dz = H / (Nz - 1);
        dt = 1 / (2 * C * (1 / dx / dx + 1 / dy / dy + 1 / dz / dz)) * 0.8;
        kx = C / (12 * dx * dx); ky = C / (12 * dy * dy); kz = C / (12 * dz * dz);
        break;
    }
    case ADV_TARGET_BURGERS3D_SGPU: {       // SingleGPU/Burgers3d_WENO5/main.cpp:53-59,130-132
        dz = H / (Nz - 1);
        C = 0.00001;
        kx = C / (12 * dx * dx); ky = C / (12 * dy * dy); kz = C / (12 * dz * dz);
        REAL dt1 = (REAL)d.coef * dx / 1.0;
        REAL dt2 = 1. / (2 * C * (1 / dx / dx + 1 / dy / dy + 1 / dz / dz)) * 0.9;
        dt = dt1 < dt2 ? dt1 : dt2;
        break;
    }
    case ADV_TARGET_BURGERS3D_MGPU: {       // MultiGPU/Burgers3d_Baseline/main.c:61-63,193
        dz = H / (Nz - 1);
        dt = (REAL)d.coef * dx / 1.0;
        break;
    }
    case ADV_TARGET_HEAT2D_MGPU: {          // MultiGPU/Diffusion2d_Baseline/main.c:57-62
        C = (REAL)d.coef;
        dt = 1 / (2 * C * (1 / dx / dx + 1 / dy / dy)) * 0.8;
        kx = C / (12 * dx * dx); ky = C / (12 * dy * dy);
        break;
    }
    case ADV_TARGET_BURGERS2D_MGPU: {       // MultiGPU/Burgers2d_Baseline/main.c:57-58,180
        dt = (REAL)d.coef * dx / 1.0;
        break;
    }
    }
    s->dx = dx; s->dy = dy; s->dz = dz; s->dt_fixed = dt; s->kx = kx; s->ky = ky; s->kz = kz; s->visc_c = C;
}

static int validate_desc(const adv_desc* d)
{
    if (!d) return fail(ADV_ERR_INVALID, "desc is NULL");
    if (d->target < ADV_TARGET_HEAT2D_SGPU || d->target > ADV_TARGET_BURGERS2D_MGPU) return fail(ADV_ERR_INVALID, "unknown target %d", d->target);
    if (d->precision != ADV_F64 && d->precision != ADV_F32) return fail(ADV_ERR_INVALID, "unknown precision %d", d->precision);
    const int minn = is_heat(d->target) ? 7 : 8;
    if (d->nx < minn || d->ny < minn) return fail(ADV_ERR_INVALID, "grid too small: %d x %d", d->nx, d->ny);
    const bool two_d = d->target == ADV_TARGET_HEAT2D_SGPU || is_2d_mgpu(d->target);
    if (!two_d && d->nz < minn) return fail(ADV_ERR_INVALID, "grid too small: nz=%d", d->nz);
    const int P = d->nranks <= 0 ? 1 : d->nranks;
    if (P > 1 && d->target == ADV_TARGET_BURGERS3D_SGPU) {
        if (d->rank < 0 || d->rank >= P) return fail(ADV_ERR_INVALID, "rank %d outside [0,%d)", d->rank, P);
        if ((d->nz - 2 * ADV_RADIUS) % P) return fail(ADV_ERR_INVALID, "nz-6=%d interior planes are not divisible by nranks=%d", d->nz - 2 * ADV_RADIUS, P);
        if ((d->nz - 2 * ADV_RADIUS) / P < 2 * ADV_RADIUS) return fail(ADV_ERR_INVALID, "slab of %d planes is thinner than 2*RADIUS", (d->nz - 2 * ADV_RADIUS) / P);
    } else if (P > 1 && !is_mgpu(d->target)) return fail(ADV_ERR_INVALID, "slab decomposition needs an MGPU target (or ADV_TARGET_BURGERS3D_SGPU)");
    if (P > 1 && is_mgpu(d->target)) {
        if (d->rank < 0 || d->rank >= P) return fail(ADV_ERR_INVALID, "rank %d outside [0,%d)", d->rank, P);
        // the reference silently drops planes (rows) when Nz %% P != 0 (main.c:64); refuse instead
        const int n = is_2d_mgpu(d->target) ? d->ny : d->nz;
        const char* ax = is_2d_mgpu(d->target) ? "ny" : "nz";
        if (n % P) return fail(ADV_ERR_INVALID, "%s=%d is not divisible by nranks=%d", ax, n, P);
        if (n / P < 2 * ADV_RADIUS) return fail(ADV_ERR_INVALID, "slab of %d %s is thinner than 2*RADIUS", n / P, is_2d_mgpu(d->target) ? "rows" : "planes");
    }
    if (!(d->coef > 0)) return fail(ADV_ERR_INVALID, "coef (C/K/CFL) must be positive");
    return ADV_OK;
}

// CUDA loads a kernel's code on its first use (lazy module loading), per device, under a process-wide lock: in the
// multi-GPU drivers (one thread per GPU) that put tens of milliseconds per rank into the first time step.  Touching
// the target's stage kernels here moves it into set-up, where the reference's drivers have their own one-time costs.
template <typename K> static void touch_kernel(K k) { cudaFuncAttributes at; (void)cudaFuncGetAttributes(&at, k); }
template <typename REAL>
static void preload_stage_kernels(int target)
{
    switch (target) {
    case ADV_TARGET_BURGERS3D_MGPU:
        touch_kernel(burgers3d_stage_pair<REAL, WENO_JS, BMODE_MGPU, 1, false>); touch_kernel(burgers3d_stage_pair<REAL, WENO_JS, BMODE_MGPU, 2, false>);
        touch_kernel(burgers3d_stage_pair<REAL, WENO_JS, BMODE_MGPU, 3, false>); touch_kernel(burgers3d_stage_pair<REAL, WENO_JS, BMODE_MGPU, 1, true>);
        touch_kernel(burgers3d_stage_pair<REAL, WENO_JS, BMODE_MGPU, 2, true>); touch_kernel(burgers3d_stage_pair<REAL, WENO_JS, BMODE_MGPU, 3, true>);
        break;
    case ADV_TARGET_BURGERS3D_SGPU:
        touch_kernel(burgers3d_stage_pair<REAL, WENO_Z, BMODE_SGPU, 1, false>); touch_kernel(burgers3d_stage_pair<REAL, WENO_Z, BMODE_SGPU, 2, false>);
        touch_kernel(burgers3d_stage_pair<REAL, WENO_Z, BMODE_SGPU, 3, false>);
        touch_kernel(burgers3d_stage_pair<REAL, WENO_Z, BMODE_SGPU, 1, true>); touch_kernel(burgers3d_stage_pair<REAL, WENO_Z, BMODE_SGPU, 2, true>);
        touch_kernel(burgers3d_stage_pair<REAL, WENO_Z, BMODE_SGPU, 3, true>);
        break;
    case ADV_TARGET_HEAT3D_MGPU: case ADV_TARGET_HEAT3D_SGPU_BASELINE: case ADV_TARGET_HEAT3D_SGPU_BLOCKING:
        touch_kernel(heat3d_stage_tma<REAL, 8, 6, 1>); touch_kernel(heat3d_stage_tma<REAL, 8, 6, 2>); touch_kernel(heat3d_stage_tma<REAL, 8, 6, 3>);
        break;
    case ADV_TARGET_BURGERS2D_MGPU:
        touch_kernel(burgers2d_stage_tile<REAL, WENO_JS, 1, false>); touch_kernel(burgers2d_stage_tile<REAL, WENO_JS, 2, false>); touch_kernel(burgers2d_stage_tile<REAL, WENO_JS, 3, false>);
        touch_kernel(burgers2d_stage_tile<REAL, WENO_JS, 1, true>); touch_kernel(burgers2d_stage_tile<REAL, WENO_JS, 2, true>); touch_kernel(burgers2d_stage_tile<REAL, WENO_JS, 3, true>);
        break;
    case ADV_TARGET_HEAT2D_MGPU:
        touch_kernel(heat2d_stage_rows<REAL, 1, false>); touch_kernel(heat2d_stage_rows<REAL, 2, false>); touch_kernel(heat2d_stage_rows<REAL, 3, false>);
        touch_kernel(heat2d_stage_rows<REAL, 1, true>); touch_kernel(heat2d_stage_rows<REAL, 2, true>); touch_kernel(heat2d_stage_rows<REAL, 3, true>);
        break;
    default: break;
    }
    (void)cudaGetLastError();
}

extern "C" int adv_create(const adv_desc* desc, adv_solver** out)
{
    if (!out) return fail(ADV_ERR_INVALID, "out is NULL");
    *out = nullptr;
    int rc = validate_desc(desc);
    if (rc) return rc;
    int ndev = 0;
    CUDA_TRY(cudaGetDeviceCount(&ndev));
    if (ndev == 0) return fail(ADV_ERR_NO_DEVICE, "no CUDA device");
    if (desc->device < 0 || desc->device >= ndev) return fail(ADV_ERR_INVALID, "device %d outside [0,%d)", desc->device, ndev);
    int major = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, desc->device));
    if (major != 10) return fail(ADV_ERR_NO_DEVICE, "device %d is sm_%d0; this library is built for sm_100a only", desc->device, major);
    CUDA_TRY(cudaSetDevice(desc->device));

    adv_solver* s = new (std::nothrow) adv_solver();
    if (!s) return fail(ADV_ERR_NOMEM, "out of host memory");
    memset((void*)s, 0, sizeof(*s));
    s->d = *desc;
    if (s->d.nranks <= 0) { s->d.nranks = 1; s->d.rank = 0; }
    if (s->d.target == ADV_TARGET_HEAT2D_SGPU || is_2d_mgpu(s->d.target)) s->d.nz = 1;
    s->esz = desc->precision == ADV_F64 ? 8 : 4;
    s->urows = is_2d_mgpu(s->d.target) ? 1 : s->d.ny;
    s->nsplit = is_2d_mgpu(s->d.target) ? s->d.ny : s->d.nz;
    if (is_sgpu_slab(s->d)) s->nsplit = s->d.nz - 2 * ADV_RADIUS;
    s->nzl = s->nsplit / s->d.nranks;
    s->nzs = (is_mgpu(desc->target) || is_sgpu_slab(s->d)) ? s->nzl + 2 * ADV_RADIUS : s->nzl;
    const size_t align = 128 / s->esz;                  // rows start on 128-byte lines
    s->pitch = ((size_t)desc->nx + align - 1) / align * align;
    s->buf_bytes = s->pitch * (size_t)s->urows * (size_t)s->nzs * s->esz;
    if (desc->precision == ADV_F64) make_constants<double>(s); else make_constants<float>(s);

#define CREATE_TRY(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { \
        int _c = fail(_e == cudaErrorMemoryAllocation ? ADV_ERR_NOMEM : ADV_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(_e)); \
        adv_destroy(s); return _c; } } while (0)
    {
        const size_t al = 512;
        size_t off = 0;
        const size_t bufb = (s->buf_bytes + al - 1) / al * al;
        for (int b = 0; b < 3; b++) { s->off_buf[b] = off; off += bufb; }
        s->off_rowacc = off;
        const size_t accb = desc->target == ADV_TARGET_BURGERS3D_MGPU ? (s->pitch * (size_t)s->nzs * s->esz + al - 1) / al * al : 0;
        off += accb;
        s->off_flags = off;
        off += 2 * al;                  // 512 B of halo / probe flags + 512 B of dynamic-CFL state
        s->arena_bytes = off;
        CREATE_TRY(cudaMalloc((void**)&s->arena, s->arena_bytes));
        for (int b = 0; b < 3; b++) s->buf[b] = s->arena + s->off_buf[b];
        if (accb) s->rowacc = s->arena + s->off_rowacc;
        CREATE_TRY(cudaMemset(s->arena + s->off_rowacc, 0, accb + 2 * al));
    }
    CREATE_TRY(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    CREATE_TRY(cudaStreamCreateWithFlags(&s->comm_stream, cudaStreamNonBlocking));
    CREATE_TRY(cudaEventCreate(&s->ev_start));
    CREATE_TRY(cudaEventCreate(&s->ev_stop));
    CREATE_TRY(cudaEventCreateWithFlags(&s->ev_boundary, cudaEventDisableTiming));
    CREATE_TRY(cudaEventCreateWithFlags(&s->ev_comm, cudaEventDisableTiming));
#undef CREATE_TRY
    if (s->d.nranks > 1) {
        // NCCL serves the ncclSend/ncclRecv halo path and the dynamic-CFL all-reduce.  A host without a loadable
        // libnccl can still run the peer-store halo path: creation succeeds with no communicator, and whatever needs
        // one later fails with "halo exchange without a communicator" (every rank sees the same missing library).
        if (!NcclApi::get(g_err, sizeof(g_err))) {
            s->comm.api = nullptr; s->comm.comm = nullptr; s->comm.nranks = s->d.nranks; s->comm.rank = s->d.rank;
        } else {
            rc = s->comm.init(s->d.nranks, s->d.rank, s->d.comm_id, g_err, sizeof(g_err));
            if (rc) { adv_destroy(s); return rc; }
        }
    }
    if (s->esz == 8) preload_stage_kernels<double>(s->d.target); else preload_stage_kernels<float>(s->d.target);
    *out = s;
    return ADV_OK;
}

extern "C" int adv_destroy(adv_solver* s)
{
    if (!s) return ADV_OK;
    cudaSetDevice(s->d.device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    if (s->comm_stream) cudaStreamSynchronize(s->comm_stream);
    s->comm.destroy();
    if (s->graph_exec) cudaGraphExecDestroy(s->graph_exec);
    for (int n = 0; n < 2; n++) if (s->peer_base[n] && s->peer_is_ipc[n]) cudaIpcCloseMemHandle(s->peer_base[n]);
    if (s->cfl_host) cudaFreeHost(s->cfl_host);
    if (s->arena) cudaFree(s->arena);
    if (s->ev_start) cudaEventDestroy(s->ev_start);
    if (s->ev_stop) cudaEventDestroy(s->ev_stop);
    if (s->ev_boundary) cudaEventDestroy(s->ev_boundary);
    if (s->ev_comm) cudaEventDestroy(s->ev_comm);
    if (s->stream) cudaStreamDestroy(s->stream);
    if (s->comm_stream) cudaStreamDestroy(s->comm_stream);
    delete s;
    return ADV_OK;
}

extern "C" size_t adv_local_elems(const adv_solver* s) { return s ? (size_t)s->d.nx * s->urows * s->nzs : 0; }
extern "C" double adv_dt(const adv_solver* s) { return s ? s->dt_fixed : 0.0; }
extern "C" long long adv_kernel_launches(const adv_solver* s) { return s ? s->launches : 0; }

extern "C" int adv_upload(adv_solver* s, const void* host)
{
    if (!s || !host) return fail(ADV_ERR_INVALID, "NULL argument");
    CUDA_TRY(cudaSetDevice(s->d.device));
    const size_t row = (size_t)s->d.nx * s->esz;
    if (s->pitch * s->esz == row)       // rows already start on 128-byte lines: one linear copy
        CUDA_TRY(cudaMemcpyAsync(s->buf[s->cur], host, row * (size_t)s->urows * s->nzs, cudaMemcpyHostToDevice, s->stream));
    else
        CUDA_TRY(cudaMemcpy2DAsync(s->buf[s->cur], s->pitch * s->esz, host, row, row, (size_t)s->urows * s->nzs,
                                   cudaMemcpyHostToDevice, s->stream));
    s->uploaded = true;
    return ADV_OK;
}

extern "C" int adv_download(adv_solver* s, void* host)
{
    if (!s || !host) return fail(ADV_ERR_INVALID, "NULL argument");
    if (!s->uploaded) return fail(ADV_ERR_STATE, "no field uploaded yet");
    CUDA_TRY(cudaSetDevice(s->d.device));
    const size_t row = (size_t)s->d.nx * s->esz;
    if (s->pitch * s->esz == row)
        CUDA_TRY(cudaMemcpyAsync(host, s->buf[s->cur], row * (size_t)s->urows * s->nzs, cudaMemcpyDeviceToHost, s->stream));
    else
        CUDA_TRY(cudaMemcpy2DAsync(host, row, s->buf[s->cur], s->pitch * s->esz, row, (size_t)s->urows * s->nzs,
                                   cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return ADV_OK;
}

extern "C" int adv_device_field(adv_solver* s, void** dptr, size_t* pitch)
{
    if (!s || !dptr) return fail(ADV_ERR_INVALID, "NULL argument");
    *dptr = s->buf[s->cur];
    if (pitch) *pitch = s->pitch;
    return ADV_OK;
}

extern "C" int adv_sync(adv_solver* s)
{
    if (!s) return fail(ADV_ERR_INVALID, "NULL solver");
    CUDA_TRY(cudaSetDevice(s->d.device));
    CUDA_TRY(cudaStreamSynchronize(s->comm_stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return ADV_OK;
}

extern "C" int adv_last_elapsed_ms(adv_solver* s, float* ms)
{
    if (!s || !ms) return fail(ADV_ERR_INVALID, "NULL argument");
    CUDA_TRY(cudaEventSynchronize(s->ev_stop));
    CUDA_TRY(cudaEventElapsedTime(ms, s->ev_start, s->ev_stop));
    return ADV_OK;
}

// ------------------------------------------------------------------------------------------
// reference initial conditions, generated on the host with libm exactly as the reference does
// ------------------------------------------------------------------------------------------
template <typename REAL>
static void reference_ic_local(const adv_solver* s, REAL* h)
{
    const adv_desc& d = s->d;
    const int nx = d.nx, ny = s->urows;                                  // y-slab targets: one row per slab unit, k = global row
    const REAL dx = (REAL)s->dx, dy = (REAL)s->dy, dz = (REAL)s->dz;
    const int NZ = is_mgpu(d.target) ? s->nsplit + 2 * ADV_RADIUS : d.nz; // global stored planes (rows)
    const int kg0 = (is_mgpu(d.target) || is_sgpu_slab(d)) ? d.rank * s->nzl : 0;            // Init_subdomain, Tools.c:181-200
    // every value depends on its own indices only: slab units are spread over host threads (same libm calls,
    // same bits; at 512^3 the serial loop was the longest part of a driver's set-up)
    const int units = s->nzs;
    int nthr = (int)std::thread::hardware_concurrency();
    if (nthr < 1) nthr = 1;
    if (nthr > 16) nthr = 16;
    if ((size_t)nx * ny * units < ((size_t)1 << 20)) nthr = 1;
    if (nthr > units) nthr = units;
    auto fill = [&](int kl0, int kl1) {
    for (int kl = kl0; kl < kl1; kl++) {
        const int k = kl + kg0;
        for (int j = 0; j < ny; j++)
            for (int i = 0; i < nx; i++) {
                const size_t o = (size_t)i + (size_t)nx * j + (size_t)nx * ny * kl;
                REAL v = 0;
                switch (d.target) {
                case ADV_TARGET_HEAT2D_SGPU:        // SingleGPU/Diffusion2d/tools.cpp:109-118
                    v = exp(-((-5 + i * dx) * (-5 + i * dx) + (-5 + j * dy) * (-5 + j * dy)) / (4 * 1.0 * 0.1));
                    if (i == 0 || i == nx - 1 || j == 0 || j == ny - 1) v = 0.0;
                    break;
                case ADV_TARGET_HEAT3D_SGPU_BASELINE:
                case ADV_TARGET_HEAT3D_SGPU_BLOCKING: // SingleGPU/Diffusion3d_baselineCode/tools.cpp:128-139
                    v = exp(-((-5 + i * dx) * (-5 + i * dx) + (-5 + j * dy) * (-5 + j * dy) + (-5 + k * dz) * (-5 + k * dz)) / (4 * 1.0 * 0.1));
                    if (i == 0 || i == nx - 1 || j == 0 || j == ny - 1 || k == 0 || k == NZ - 1) v = 0.0;
                    break;
                case ADV_TARGET_HEAT3D_MGPU:        // MultiGPU/Diffusion3d_Baseline/Tools.c:129-143 (case 1)
                    v = (i >= nx / 4 && i < 3 * nx / 4 && j >= ny / 4 && j < 3 * ny / 4 && k >= NZ / 4 && k < 3 * NZ / 4) ? 1. : 0.;
                    break;
                case ADV_TARGET_BURGERS3D_SGPU: {   // SingleGPU/Burgers3d_WENO5/tools.cpp:126-136
                    const REAL x = i * dx, y = j * dy, z = k * dz;
                    v = 1.0 * exp(-((x - 1.0) * (x - 1.0) + (y - 1.0) * (y - 1.0) + (z - 1.0) * (z - 1.0)) / 0.1);
                    break;
                }
                case ADV_TARGET_BURGERS3D_MGPU:     // MultiGPU/Burgers3d_Baseline/Tools.c:158-173 (macro pastes
                                                    // unparenthesised arguments: ((a*d)*a)*d)
                    if (i == 0 || i == nx - 1 || j == 0 || j == ny - 1 || k == 0 || k == NZ - 1) v = 0.0;
                    else v = 1.0 * exp(-(((0.5 * (nx - 1) - i) * dx * (0.5 * (nx - 1) - i) * dx) +
                                          ((0.5 * (ny - 1) - j) * dy * (0.5 * (ny - 1) - j) * dy) +
                                          ((0.5 * (NZ - 1) - k) * dz * (0.5 * (NZ - 1) - k) * dz)) / 0.1);
                    break;
                case ADV_TARGET_HEAT2D_MGPU:        // MultiGPU/Diffusion2d_Baseline/Tools.c:108-121 (radius rounded by sqrtf)
                    v = (sqrtf((float)((i * 1. / (nx - 1) - 0.5) * (i * 1. / (nx - 1) - 0.5) +
                                       (k * 1. / (NZ - 1) - 0.5) * (k * 1. / (NZ - 1) - 0.5))) < 0.2) ? 1.0 : 0.0;
                    break;
                case ADV_TARGET_BURGERS2D_MGPU:     // MultiGPU/Burgers2d_Baseline/Tools.c:108-121, BurgersMPICUDA.h:60
                    if (i == 0 || i == nx - 1 || k == 0 || k == NZ - 1) v = 0.0;
                    else v = 1.0 * exp(-(((0.5 * (nx - 1) - i) * dx * (0.5 * (nx - 1) - i) * dx) +
                                          ((0.5 * (NZ - 1) - k) * dy * (0.5 * (NZ - 1) - k) * dy)) / 0.1);
                    break;
                }
                h[o] = v;
            }
    }
    };
    if (nthr == 1) { fill(0, units); return; }
    std::vector<std::thread> pool;
    for (int t = 0; t < nthr; t++)
        pool.emplace_back(fill, (int)((long long)units * t / nthr), (int)((long long)units * (t + 1) / nthr));
    for (auto& th : pool) th.join();
}

extern "C" int adv_init_reference_ic(adv_solver* s)
{
    if (!s) return fail(ADV_ERR_INVALID, "NULL solver");
    const size_t n = adv_local_elems(s);
    void* h = malloc(n * s->esz);
    if (!h) return fail(ADV_ERR_NOMEM, "out of host memory");
    if (s->esz == 8) reference_ic_local<double>(s, (double*)h); else reference_ic_local<float>(s, (float*)h);
    int rc = adv_upload(s, h);
    if (!rc) { cudaError_t e = cudaStreamSynchronize(s->stream); if (e != cudaSuccess) rc = fail(ADV_ERR_CUDA, "upload: %s", cudaGetErrorString(e)); }
    free(h);
    return rc;
}

// ------------------------------------------------------------------------------------------
// The same initial conditions generated on the device, straight into the slab: at 1024^3 the
// reference fills the whole padded domain serially on every rank (main.c:70-73, Tools.c:124-176).  For
// throughput runs only: the device's exp() is not the host libm's, so the values agree with
// adv_init_reference_ic to the last bit or two, not bit for bit -- parity runs keep the host generator.
// ------------------------------------------------------------------------------------------
template <typename REAL>
__global__ void __launch_bounds__(256)
reference_ic_kernel(REAL* __restrict__ h, int target, int nx, int ny, int units, long long pitch, int NZ, int kg0,
                    REAL dx, REAL dy, REAL dz)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= nx) return;
    for (int kl = blockIdx.z; kl < units; kl += gridDim.z) {
        const int k = kl + kg0;
        REAL v = 0;
        switch (target) {
        case ADV_TARGET_HEAT2D_SGPU:
            v = exp(-((-5 + i * dx) * (-5 + i * dx) + (-5 + j * dy) * (-5 + j * dy)) / (4 * 1.0 * 0.1));
            if (i == 0 || i == nx - 1 || j == 0 || j == ny - 1) v = 0.0;
            break;
        case ADV_TARGET_HEAT3D_SGPU_BASELINE:
        case ADV_TARGET_HEAT3D_SGPU_BLOCKING:
            v = exp(-((-5 + i * dx) * (-5 + i * dx) + (-5 + j * dy) * (-5 + j * dy) + (-5 + k * dz) * (-5 + k * dz)) / (4 * 1.0 * 0.1));
            if (i == 0 || i == nx - 1 || j == 0 || j == ny - 1 || k == 0 || k == NZ - 1) v = 0.0;
            break;
        case ADV_TARGET_HEAT3D_MGPU:
            v = (i >= nx / 4 && i < 3 * nx / 4 && j >= ny / 4 && j < 3 * ny / 4 && k >= NZ / 4 && k < 3 * NZ / 4) ? 1. : 0.;
            break;
        case ADV_TARGET_BURGERS3D_SGPU: {
            const REAL x = i * dx, y = j * dy, z = k * dz;
            v = 1.0 * exp(-((x - 1.0) * (x - 1.0) + (y - 1.0) * (y - 1.0) + (z - 1.0) * (z - 1.0)) / 0.1);
            break;
        }
        case ADV_TARGET_BURGERS3D_MGPU:
            if (i == 0 || i == nx - 1 || j == 0 || j == ny - 1 || k == 0 || k == NZ - 1) v = 0.0;
            else v = 1.0 * exp(-(((0.5 * (nx - 1) - i) * dx * (0.5 * (nx - 1) - i) * dx) +
                                  ((0.5 * (ny - 1) - j) * dy * (0.5 * (ny - 1) - j) * dy) +
                                  ((0.5 * (NZ - 1) - k) * dz * (0.5 * (NZ - 1) - k) * dz)) / 0.1);
            break;
        case ADV_TARGET_HEAT2D_MGPU:        // y-slab: one row per unit, k = global row, ny == 1
            v = (sqrtf((float)((i * 1. / (nx - 1) - 0.5) * (i * 1. / (nx - 1) - 0.5) +
                               (k * 1. / (NZ - 1) - 0.5) * (k * 1. / (NZ - 1) - 0.5))) < 0.2) ? 1.0 : 0.0;
            break;
        case ADV_TARGET_BURGERS2D_MGPU:
            if (i == 0 || i == nx - 1 || k == 0 || k == NZ - 1) v = 0.0;
            else v = 1.0 * exp(-(((0.5 * (nx - 1) - i) * dx * (0.5 * (nx - 1) - i) * dx) +
                                  ((0.5 * (NZ - 1) - k) * dy * (0.5 * (NZ - 1) - k) * dy)) / 0.1);
            break;
        }
        h[(long long)i + pitch * ((long long)j + (long long)ny * kl)] = v;
    }
}

extern "C" int adv_init_reference_ic_device(adv_solver* s)
{
    if (!s) return fail(ADV_ERR_INVALID, "NULL solver");
    CUDA_TRY(cudaSetDevice(s->d.device));
    const adv_desc& d = s->d;
    const int NZ = is_mgpu(d.target) ? s->nsplit + 2 * ADV_RADIUS : d.nz;
    const int kg0 = (is_mgpu(d.target) || is_sgpu_slab(d)) ? d.rank * s->nzl : 0;
    dim3 block(256, 1, 1), grid((d.nx + 255) / 256, s->urows, s->nzs < 1024 ? s->nzs : 1024);
    if (s->esz == 8)
        reference_ic_kernel<double><<<grid, block, 0, s->stream>>>((double*)s->buf[s->cur], d.target, d.nx, s->urows, s->nzs,
                                                                      (long long)s->pitch, NZ, kg0, s->dx, s->dy, s->dz);
    else
        reference_ic_kernel<float><<<grid, block, 0, s->stream>>>((float*)s->buf[s->cur], d.target, d.nx, s->urows, s->nzs,
                                                                     (long long)s->pitch, NZ, kg0, (float)s->dx, (float)s->dy, (float)s->dz);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    s->uploaded = true;
    return ADV_OK;
}

// ------------------------------------------------------------------------------------------
// one time step = three fused stages over ping-pong buffers A (u^n), B, C:
//   B = S1(A);  C = S2(B; A);  B = S3(C; A);  then A <-> B.      (no uo copy, no Lu array)
// With ranks > 1 each stage computes the two 3-plane boundary slabs first, hands them to the
// halo exchange on the comm stream, and computes the interior while they travel.
// ------------------------------------------------------------------------------------------
static int launch_stage_peer(adv_solver* s, int stage, double dt, const void* in, const void* uo, void* out, int kb, int ke,
                             cudaStream_t st, PeerOut peer)
{
    if (kb >= ke && !peer.fused) return ADV_OK;
    if (kb > ke) ke = kb;
    adv_stage_args a;
    memset(&a, 0, sizeof(a));
    a.target = s->d.target; a.flags = s->d.flags; a.stage = stage;
    const bool yslab = is_2d_mgpu(s->d.target);
    a.nx = s->d.nx; a.ny = yslab ? s->nzs : s->d.ny; a.nzs = yslab ? 1 : s->nzs; a.pitch = s->pitch;
    a.k_begin = kb; a.k_end = ke;
    a.own_lo = is_mgpu(s->d.target) ? ADV_RADIUS : 3;
    a.own_hi = s->nzs - 3;
    a.k_global0 = (yslab || is_sgpu_slab(s->d)) ? s->d.rank * s->nzl : 0;
    a.dt = dt; a.dx = s->dx; a.dy = s->dy; a.dz = s->dz; a.kx = s->kx; a.ky = s->ky; a.kz = s->kz;
    int rc;
    const bool f64 = s->esz == 8;
    switch (s->d.target) {
    case ADV_TARGET_HEAT2D_SGPU:
    case ADV_TARGET_HEAT2D_MGPU:
        rc = f64 ? heat2d_stage_impl<double>(&a, (const double*)in, (const double*)uo, (double*)out, st, peer)
                 : heat2d_stage_impl<float>(&a, (const float*)in, (const float*)uo, (float*)out, st, peer);
        break;
    case ADV_TARGET_BURGERS2D_MGPU:
        rc = f64 ? burgers2d_stage_impl<double>(&a, (const double*)in, (const double*)uo, (double*)out, st, peer)
                 : burgers2d_stage_impl<float>(&a, (const float*)in, (const float*)uo, (float*)out, st, peer);
        break;
    case ADV_TARGET_BURGERS3D_SGPU:
    case ADV_TARGET_BURGERS3D_MGPU:
        if (s->cfl_on) {
            CflDevState* cs = (CflDevState*)(s->arena + s->off_flags + 512);
            peer.cfl_dt = &cs->dt; peer.cfl_done = &cs->done; peer.cfl_umax = &cs->local_max;
        }
        rc = f64 ? burgers3d_stage_impl<double>(&a, (const double*)in, (const double*)uo, (double*)out, (double*)s->rowacc, st, peer)
                 : burgers3d_stage_impl<float>(&a, (const float*)in, (const float*)uo, (float*)out, (float*)s->rowacc, st, peer);
        break;
    default:
        rc = f64 ? heat3d_stage_impl<double>(&a, (const double*)in, (const double*)uo, (double*)out, st, peer)
                 : heat3d_stage_impl<float>(&a, (const float*)in, (const float*)uo, (float*)out, st, peer);
    }
    if (!rc) s->launches++;
    return rc;
}

static int launch_stage(adv_solver* s, int stage, double dt, const void* in, const void* uo, void* out, int kb, int ke, cudaStream_t st)
{
    return launch_stage_peer(s, stage, dt, in, uo, out, kb, ke, st, kNoPeer);
}

// ---- fused halo exchange over peer memory --------------------------------------------------------
// flag words inside every arena: [0] written by the left neighbour, [32] by the right one (128 B apart)
enum { FLAG_FROM_LEFT = 0, FLAG_FROM_RIGHT = 32 };

__global__ void halo_signal_kernel(volatile unsigned* flag, unsigned value)
{
    __threadfence_system();         // the boundary-slab kernel before us in the stream has completed its peer stores
    *flag = value;
    __threadfence_system();
}
__global__ void halo_wait_kernel(const volatile unsigned* flag, unsigned value)
{
    while ((int)(*flag - value) < 0) __nanosleep(200);
}

typedef CUresult (*stream_wait32_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
static stream_wait32_fn lookup_stream_wait32()
{
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (!getenv("ADV_HALO_SPIN_WAIT") &&
        cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
        return (stream_wait32_fn)p;
    return nullptr;
}
static stream_wait32_fn get_stream_wait32()
{
    static const stream_wait32_fn fn = lookup_stream_wait32();     // thread-safe one-time lookup
    return fn;
}

static int halo_wait(adv_solver* s, int which, unsigned value, cudaStream_t st = nullptr)
{
    if (!st) st = s->stream;
    unsigned* flag = (unsigned*)(s->arena + s->off_flags) + which;
    stream_wait32_fn w = get_stream_wait32();
    if (w) {
        CUresult r = w((CUstream)st, (CUdeviceptr)flag, value, 0 /* CU_STREAM_WAIT_VALUE_GEQ */);
        if (r == CUDA_SUCCESS) return ADV_OK;
    }
    halo_wait_kernel<<<1, 1, 0, st>>>(flag, value);
    CUDA_TRY(cudaGetLastError());
    s->launches++;
    return ADV_OK;
}

static int halo_signal(adv_solver* s, int nbr, int which, unsigned value, cudaStream_t st = nullptr)
{
    if (!st) st = s->stream;
    unsigned* flag = (unsigned*)(s->peer_base[nbr] + s->off_flags) + which;
    halo_signal_kernel<<<1, 1, 0, st>>>(flag, value);
    CUDA_TRY(cudaGetLastError());
    s->launches++;
    return ADV_OK;
}

static int launch_stage_peer(adv_solver* s, int stage, double dt, const void* in, const void* uo, void* out, int kb, int ke,
                             cudaStream_t st, PeerOut peer);

// One stage with the exchange fused into the boundary-slab kernels: they store their three planes both
// locally and, through NVLink peer mappings, into the neighbour's ghost planes; a flag per neighbour
// says "stage c has landed".  A rank starts stage c once both neighbours have signalled c-1, which is
// also when they have stopped reading the ghost planes this stage overwrites.  No NCCL, no comm stream.
// 3-D targets whose stage kernel takes a SlabSched: the whole stage is ONE launch (boundary windows first, flag wait
// and flag release inside the kernel) instead of two stream waits + two 3-plane launches + two signal kernels + the
// interior launch.  ADV_HALO_MULTILAUNCH=1 keeps the old schedule (A/B).
static bool stage_fusable(const adv_solver* s)
{
    if (env_on("ADV_HALO_MULTILAUNCH") || !get_tmap_encode()) return false;
    // heat: measured on 2 B200s (tools/r02_ab_fused.sh, profiles/r02_ab_fused_2gpu.log) the one-launch stage is
    // 0.5 % (512^3 per GPU) to 4.5 % (1024^3 strong) SLOWER than boundary launches + interior launch -- the short
    // boundary CTAs of the bandwidth-bound kernel do better with the GPU to themselves -- so it is opt-in there
    if (s->d.target == ADV_TARGET_HEAT3D_MGPU)
        return env_on("ADV_HEAT_ONE_LAUNCH") && !env_on("ADV_HEAT_DIRECT") && s->d.nx % 2 == 0 && s->d.nx >= 8;
    if (s->d.target == ADV_TARGET_BURGERS3D_MGPU || s->d.target == ADV_TARGET_BURGERS3D_SGPU)
        return !env_on("ADV_BURGERS_DIRECT") && !env_on("ADV_BURGERS_RING") && s->d.nx >= 8;
    return false;
}

enum { STAGE_TICKET_LEFT = 88, STAGE_TICKET_RIGHT = 92 };

static int do_stage_p2p(adv_solver* s, int stage, double dt, const void* in, const void* uo, void* out)
{
    const int P = s->d.nranks, r = s->d.rank, R = ADV_RADIUS;
    const unsigned c = ++s->stage_counter;
    int ob = -1;
    for (int b = 0; b < 3; b++) if (s->buf[b] == out) ob = b;
    if (ob < 0) return fail(ADV_ERR_STATE, "stage output is not one of the solver's buffers");
    const long long plane = (long long)s->pitch * s->urows;
    int rc;
    // heat, A/B only (ADV_HEAT_TWO_LAUNCH=1): both boundary windows with in-kernel flag wait/release in ONE launch, then
    // the interior launch.  Measured on 4 B200s (profiles/r02_ab_fused_4gpu.log), 512^3 per GPU: round-1 schedule
    // (2 stream waits + 2 boundary launches + 2 signal kernels + interior) 1.521 ms per step, everything in one launch
    // 1.544, this variant 1.575 -- in-kernel flag waits cost the bandwidth-bound kernel more than the launches they
    // save, so heat keeps the round-1 schedule; Burgers (7.50 -> 7.44 ms) takes the one-launch stage.
    const bool two_launch = !stage_fusable(s) && s->d.target == ADV_TARGET_HEAT3D_MGPU && env_on("ADV_HEAT_TWO_LAUNCH") &&
                            !env_on("ADV_HEAT_DIRECT") && get_tmap_encode() && s->d.nx % 2 == 0 && s->d.nx >= 8;
    if (stage_fusable(s) || two_launch) {
        unsigned* flags = (unsigned*)(s->arena + s->off_flags);
        SlabSched f;
        memset(&f, 0, sizeof(f));
        f.wait_val = c - 1; f.signal_val = c;
        if (r > 0) {            // my planes [3,6) are the left neighbour's ghost planes [nzs-3, nzs)
            const int w = f.nb++;
            f.wb[w] = R; f.we[w] = 2 * R;
            f.out2[w] = s->peer_base[0] + s->off_buf[ob]; f.off2[w] = (long long)(s->nzs - 2 * R) * plane;
            f.wait_flag[w] = flags + FLAG_FROM_LEFT;
            f.peer_flag[w] = (unsigned*)(s->peer_base[0] + s->off_flags) + FLAG_FROM_RIGHT;
            f.ticket[w] = flags + STAGE_TICKET_LEFT;
        }
        if (r < P - 1) {        // my planes [nzs-6, nzs-3) are the right neighbour's ghost planes [0, 3)
            const int w = f.nb++;
            f.wb[w] = s->nzs - 2 * R; f.we[w] = s->nzs - R;
            f.out2[w] = s->peer_base[1] + s->off_buf[ob]; f.off2[w] = -(long long)(s->nzs - 2 * R) * plane;
            f.wait_flag[w] = flags + FLAG_FROM_RIGHT;
            f.peer_flag[w] = (unsigned*)(s->peer_base[1] + s->off_flags) + FLAG_FROM_LEFT;
            f.ticket[w] = flags + STAGE_TICKET_RIGHT;
        }
        PeerOut po = kNoPeer;
        po.fused = &f;
        const int flo = (r > 0) ? 2 * R : 0, fhi = (r < P - 1) ? s->nzs - 2 * R : s->nzs;
        if (!two_launch) return launch_stage_peer(s, stage, dt, in, uo, out, flo, fhi, s->stream, po);
        if ((rc = launch_stage_peer(s, stage, dt, in, uo, out, flo, flo, s->stream, po))) return rc;     // boundary windows only
        return launch_stage_peer(s, stage, dt, in, uo, out, flo, fhi, s->stream, kNoPeer);
    }
    // Multi-launch schedule.  For heat 3-D the boundary launches, their flag waits and signals go to the side stream,
    // forked from the main stream at the start of the stage and joined at the start of the next one, so that the
    // interior launch fills the SMs the short boundary CTAs leave idle: 4 B200s, 512^3 per GPU, 1.484 -> 1.426 ms per
    // step (91.8 -> 95.6 % of the one-GPU rate), 1024^3 strong 3.06 -> 2.99 ms (profiles/r02_ab_side_4gpu.log).
    // ADV_HALO_ONE_STREAM=1 keeps everything on one stream; ADV_HALO_SIDE_STREAM=1 turns the side stream on for the
    // other multi-launch targets too (parity-tested, not timed).
    const bool side = !env_on("ADV_HALO_ONE_STREAM") && (s->d.target == ADV_TARGET_HEAT3D_MGPU || env_on("ADV_HALO_SIDE_STREAM"));
    cudaStream_t bs = side ? s->comm_stream : s->stream;
    if (side) {
        CUDA_TRY(cudaStreamWaitEvent(s->stream, s->ev_comm, 0));        // join the previous stage's boundary work
        CUDA_TRY(cudaEventRecord(s->ev_boundary, s->stream));           // fork: everything up to the previous interior launch
        CUDA_TRY(cudaStreamWaitEvent(bs, s->ev_boundary, 0));
    }
    if (r > 0 && (rc = halo_wait(s, FLAG_FROM_LEFT, c - 1, bs))) return rc;
    if (r < P - 1 && (rc = halo_wait(s, FLAG_FROM_RIGHT, c - 1, bs))) return rc;
    if (r > 0) {            // my planes [3,6) are the left neighbour's ghost planes [nzs-3, nzs)
        PeerOut po = {s->peer_base[0] + s->off_buf[ob], (long long)(s->nzs - 2 * R) * plane};
        if ((rc = launch_stage_peer(s, stage, dt, in, uo, out, R, 2 * R, bs, po))) return rc;
    }
    if (r < P - 1) {        // my planes [nzs-6, nzs-3) are the right neighbour's ghost planes [0, 3)
        PeerOut po = {s->peer_base[1] + s->off_buf[ob], -(long long)(s->nzs - 2 * R) * plane};
        if ((rc = launch_stage_peer(s, stage, dt, in, uo, out, s->nzs - 2 * R, s->nzs - R, bs, po))) return rc;
    }
    if (r > 0 && (rc = halo_signal(s, 0, FLAG_FROM_RIGHT, c, bs))) return rc;
    if (r < P - 1 && (rc = halo_signal(s, 1, FLAG_FROM_LEFT, c, bs))) return rc;
    if (side) CUDA_TRY(cudaEventRecord(s->ev_comm, bs));
    const int ilo = (r > 0) ? 2 * R : 0;
    const int ihi = (r < P - 1) ? s->nzs - 2 * R : s->nzs;
    return launch_stage_peer(s, stage, dt, in, uo, out, ilo, ihi, s->stream, kNoPeer);
}

static int do_stage(adv_solver* s, int stage, double dt, const void* in, const void* uo, void* out)
{
    if (s->d.nranks > 1 && s->p2p) return do_stage_p2p(s, stage, dt, in, uo, out);
    const int P = s->d.nranks, r = s->d.rank;
    if (P == 1) return launch_stage(s, stage, dt, in, uo, out, 0, s->nzs, s->stream);

    // planes this rank produces itself: owned [3, nzs-3) plus the global pad planes at the two ends
    const int lo = (r == 0) ? 0 : ADV_RADIUS;
    const int hi = (r == P - 1) ? s->nzs : s->nzs - ADV_RADIUS;
    const int R = ADV_RADIUS;
    int rc;
    // the comm stream of the previous stage must be done with `in` ghosts before we read them
    CUDA_TRY(cudaStreamWaitEvent(s->stream, s->ev_comm, 0));
    // boundary slabs first: [3,6) goes to the left neighbour, [nzs-6,nzs-3) to the right one
    if (r > 0)     { rc = launch_stage(s, stage, dt, in, uo, out, R, 2 * R, s->stream); if (rc) return rc; }
    if (r < P - 1) { rc = launch_stage(s, stage, dt, in, uo, out, s->nzs - 2 * R, s->nzs - R, s->stream); if (rc) return rc; }
    CUDA_TRY(cudaEventRecord(s->ev_boundary, s->stream));
    CUDA_TRY(cudaStreamWaitEvent(s->comm_stream, s->ev_boundary, 0));
    rc = s->comm.exchange(out, s->pitch * (size_t)s->urows, s->nzs, s->esz, s->comm_stream, g_err, sizeof(g_err));
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(s->ev_comm, s->comm_stream));
    // interior (and the global-end pad planes) while the slabs travel
    const int ilo = (r > 0) ? 2 * R : lo;
    const int ihi = (r < P - 1) ? s->nzs - 2 * R : hi;
    return launch_stage(s, stage, dt, in, uo, out, ilo, ihi, s->stream);
}

// End of a run of steps on a multi-rank solver: make "this rank's stream is idle" imply "this rank's ghost
// planes hold the neighbours' last stage".  NCCL path: wait for the comm stream's last exchange.  Peer-store
// path: the neighbours write our ghost planes themselves, so wait until both have signalled the last stage
// executed -- otherwise adv_download could return stale ghost planes and an adv_upload that follows could
// overwrite freshly stored ones (the next step's flag wait would then pass on wrong data).
static int halo_quiesce(adv_solver* s)
{
    if (s->d.nranks <= 1) return ADV_OK;
    if (!s->p2p) { CUDA_TRY(cudaStreamWaitEvent(s->stream, s->ev_comm, 0)); return ADV_OK; }
    int rc;
    if (s->stage_counter == 0) return ADV_OK;
    CUDA_TRY(cudaStreamWaitEvent(s->stream, s->ev_comm, 0));        // side-stream boundary work of the last stage, if any
    if (s->d.rank > 0 && (rc = halo_wait(s, FLAG_FROM_LEFT, s->stage_counter))) return rc;
    if (s->d.rank < s->d.nranks - 1 && (rc = halo_wait(s, FLAG_FROM_RIGHT, s->stage_counter))) return rc;
    return ADV_OK;
}

// ---- the exchange alone, timed (SURVEY 8d "Halo metric") -------------------------------------------
// What a stage ships -- three boundary planes per neighbour -- and nothing else: per repetition every rank
// stores its boundary planes into its neighbours' ghost planes (the values already there: the field does not
// change), the last CTA of each copy raises the neighbour's probe flag, and the rank waits for its own probe
// flags.  CUDA events around the loop give the time of one complete exchange as this rank sees it, neighbours'
// sends included.  NCCL mode times HaloComm::exchange the same way.  Probe flags are separate words of the
// flag block, so the stage counters of the time loop are untouched.
enum { FLAG_PROBE_FROM_LEFT = 96, FLAG_PROBE_FROM_RIGHT = 112, PROBE_TICKET_LEFT = 80, PROBE_TICKET_RIGHT = 84 };

__global__ void __launch_bounds__(256)
halo_probe_copy_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n16, unsigned* ticket,
                       volatile unsigned* peer_flag, unsigned value)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(ticket, 1u);
        if (t == gridDim.x - 1) {           // last CTA: every other CTA's stores are fenced before its ticket
            *ticket = 0;
            __threadfence_system();
            *peer_flag = value;
        }
    }
}

extern "C" int adv_halo_probe(adv_solver* s, int reps, float* ms_per_exchange, double* bytes_sent_per_exchange)
{
    if (!s || !ms_per_exchange) return fail(ADV_ERR_INVALID, "NULL argument");
    if (s->d.nranks < 2) return fail(ADV_ERR_INVALID, "the halo probe needs nranks > 1");
    if (!s->uploaded) return fail(ADV_ERR_STATE, "adv_halo_probe before adv_upload");
    if (reps < 1) reps = 1;
    CUDA_TRY(cudaSetDevice(s->d.device));
    const int P = s->d.nranks, r = s->d.rank, R = ADV_RADIUS;
    const size_t plane_b = s->pitch * (size_t)s->urows * s->esz, bytes = (size_t)R * plane_b;
    const int nnb = (r > 0 ? 1 : 0) + (r < P - 1 ? 1 : 0);
    if (bytes_sent_per_exchange) *bytes_sent_per_exchange = (double)bytes * nnb;
    char* field = (char*)s->buf[s->cur];
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    int rc = ADV_OK;
    if (!s->p2p) {
        CUDA_TRY(cudaStreamWaitEvent(s->comm_stream, s->ev_comm, 0));
        CUDA_TRY(cudaEventRecord(e0, s->comm_stream));
        for (int n = 0; n < reps && !rc; n++)
            rc = s->comm.exchange(field, s->pitch * (size_t)s->urows, s->nzs, s->esz, s->comm_stream, g_err, sizeof(g_err));
        CUDA_TRY(cudaEventRecord(e1, s->comm_stream));
        CUDA_TRY(cudaEventRecord(s->ev_comm, s->comm_stream));
    } else {
        stream_wait32_fn w = get_stream_wait32();
        unsigned* flags = (unsigned*)(s->arena + s->off_flags);
        const size_t off_cur = (size_t)(field - s->arena);
        int dev_sms = 148;
        cudaDeviceGetAttribute(&dev_sms, cudaDevAttrMultiProcessorCount, s->d.device);
        CUDA_TRY(cudaEventRecord(e0, s->stream));
        for (int n = 0; n < reps; n++) {
            const unsigned c = ++s->probe_counter;
            if (r > 0)        // my planes [3,6) -> left neighbour's ghost planes [nzs-3, nzs)
                halo_probe_copy_kernel<<<8 * dev_sms, 256, 0, s->stream>>>(
                    (const uint4*)(field + (size_t)R * plane_b), (uint4*)(s->peer_base[0] + off_cur + (size_t)(s->nzs - R) * plane_b), bytes / 16,
                    flags + PROBE_TICKET_LEFT, (volatile unsigned*)(s->peer_base[0] + s->off_flags) + FLAG_PROBE_FROM_RIGHT, c);
            if (r < P - 1)    // my planes [nzs-6, nzs-3) -> right neighbour's ghost planes [0, 3)
                halo_probe_copy_kernel<<<8 * dev_sms, 256, 0, s->stream>>>(
                    (const uint4*)(field + (size_t)(s->nzs - 2 * R) * plane_b), (uint4*)(s->peer_base[1] + off_cur), bytes / 16,
                    flags + PROBE_TICKET_RIGHT, (volatile unsigned*)(s->peer_base[1] + s->off_flags) + FLAG_PROBE_FROM_LEFT, c);
            CUDA_TRY(cudaGetLastError());
            const int which[2] = {FLAG_PROBE_FROM_LEFT, FLAG_PROBE_FROM_RIGHT};
            for (int side = 0; side < 2; side++) {
                if ((side == 0 && r == 0) || (side == 1 && r == P - 1)) continue;
                bool done = false;
                if (w) done = w((CUstream)s->stream, (CUdeviceptr)(flags + which[side]), c, 0) == CUDA_SUCCESS;
                if (!done) { halo_wait_kernel<<<1, 1, 0, s->stream>>>(flags + which[side], c); CUDA_TRY(cudaGetLastError()); }
            }
        }
        CUDA_TRY(cudaEventRecord(e1, s->stream));
    }
    if (!rc) {
        CUDA_TRY(cudaEventSynchronize(e1));
        float ms = 0;
        CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
        *ms_per_exchange = ms / reps;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return rc;
}

static int do_step(adv_solver* s, double dt)
{
    void* A = s->buf[s->cur];
    void* B = s->buf[(s->cur + 1) % 3];
    void* C = s->buf[(s->cur + 2) % 3];
    int rc;
    if ((rc = do_stage(s, 1, dt, A, A, B))) return rc;
    if ((rc = do_stage(s, 2, dt, B, A, C))) return rc;
    if ((rc = do_stage(s, 3, dt, C, A, B))) return rc;
    s->cur = (s->cur + 1) % 3;
    return ADV_OK;
}

extern "C" int adv_step(adv_solver* s, int nsteps)
{
    if (!s) return fail(ADV_ERR_INVALID, "NULL solver");
    if (!s->uploaded) return fail(ADV_ERR_STATE, "adv_step before adv_upload");
    if (nsteps < 0) return fail(ADV_ERR_INVALID, "nsteps < 0");
    CUDA_TRY(cudaSetDevice(s->d.device));
    const bool no_graph = env_on("ADV_NO_GRAPH");
    const size_t cells = (size_t)s->d.nx * s->urows * s->nzs;
    const bool use_graph = !no_graph && s->d.nranks == 1 && nsteps >= 12 && cells <= ((size_t)1 << 25);   // up to 256^3 (65 us heat stages)
    if (use_graph && !s->graph_exec) {
        // capture three steps starting from the current buffer assignment
        cudaGraph_t g = nullptr;
        const long long l0 = s->launches;
        const int cur0 = s->cur;
        CUDA_TRY(cudaStreamBeginCapture(s->stream, cudaStreamCaptureModeThreadLocal));
        int rc = ADV_OK;
        for (int n = 0; n < 3 && !rc; n++) rc = do_step(s, s->dt_fixed);
        cudaError_t e = cudaStreamEndCapture(s->stream, &g);
        s->launches = l0;
        s->cur = cur0;
        if (rc) { if (g) cudaGraphDestroy(g); return rc; }
        if (e != cudaSuccess) return fail(ADV_ERR_CUDA, "cudaStreamEndCapture: %s", cudaGetErrorString(e));
        e = cudaGraphInstantiate(&s->graph_exec, g, 0);
        cudaGraphDestroy(g);
        if (e != cudaSuccess) { s->graph_exec = nullptr; return fail(ADV_ERR_CUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(e)); }
        s->graph_cur = cur0;
    }
    CUDA_TRY(cudaEventRecord(s->ev_start, s->stream));
    int n = 0;
    if (use_graph) {
        while (n < nsteps && s->cur != s->graph_cur) { int rc = do_step(s, s->dt_fixed); if (rc) return rc; n++; }
        for (; n + 3 <= nsteps; n += 3) { CUDA_TRY(cudaGraphLaunch(s->graph_exec, s->stream)); s->launches += 9; }
    }
    for (; n < nsteps; n++) {
        int rc = do_step(s, s->dt_fixed);
        if (rc) return rc;
    }
    { int rc = halo_quiesce(s); if (rc) return rc; }
    CUDA_TRY(cudaEventRecord(s->ev_stop, s->stream));
    return ADV_OK;
}

template <typename REAL>
static int run_until_impl(adv_solver* s, double tEnd_d, int max_steps, double* dt_last, double* steps, double* t_reached)
{
    // main.c:190-196 / main.cpp:127-136: REAL t, it, dt; fixed dt, clipped on the last step (Burgers).
    // MultiGPU heat (main.c:189-193) does not clip.
    const bool clip = !is_heat(s->d.target);
    const REAL tEnd = (REAL)tEnd_d;
    REAL dt = 0, it = 0, t = 0;
    while (t < tEnd && (max_steps < 0 || it < max_steps)) {
        dt = (REAL)s->dt_fixed;
        if (clip && (t + dt) > tEnd) { dt = tEnd - t; }
        t += dt; it += 1;
        int rc = do_step(s, (double)dt);
        if (rc) return rc;
    }
    if (dt_last) *dt_last = dt;
    if (steps) *steps = it;
    if (t_reached) *t_reached = t;
    return ADV_OK;
}

extern "C" int adv_run_until(adv_solver* s, double tEnd, int max_steps, double* dt_last, double* steps, double* t_reached)
{
    if (!s) return fail(ADV_ERR_INVALID, "NULL solver");
    if (!s->uploaded) return fail(ADV_ERR_STATE, "adv_run_until before adv_upload");
    CUDA_TRY(cudaSetDevice(s->d.device));
    CUDA_TRY(cudaEventRecord(s->ev_start, s->stream));
    int rc = s->esz == 8 ? run_until_impl<double>(s, tEnd, max_steps, dt_last, steps, t_reached)
                         : run_until_impl<float>(s, tEnd, max_steps, dt_last, steps, t_reached);
    if (rc) return rc;
    if ((rc = halo_quiesce(s))) return rc;
    CUDA_TRY(cudaEventRecord(s->ev_stop, s->stream));
    return ADV_OK;
}

// ------------------------------------------------------------------------------------------
// the prototype's dynamic time step.  dt = CFL*dx/max|u| is recomputed before every step from one
// reduction pass over the current field (LFWENO5FDM3d.m:71); the CUDA drivers of the reference fix max|u| = 1
// (main.c:193), which adv_run_until reproduces.  The 8-byte result crosses to the host once per step.
// ------------------------------------------------------------------------------------------
template <typename REAL>
static int run_until_cfl_impl(adv_solver* s, double tEnd_d, int max_steps, double* dt_last, double* steps, double* t_reached)
{
    unsigned long long* d_max = (unsigned long long*)(s->arena + s->off_flags + 256);      // spare words of the flag block
    const REAL tEnd = (REAL)tEnd_d, CFL = (REAL)s->d.coef, dx = (REAL)s->dx;
    const long long rows = (long long)s->urows * s->nzs;
    const int grid = (int)(rows < current_sm_count() * 8 ? rows : current_sm_count() * 8);
    REAL dt = 0, it = 0, t = 0;
    while (t < tEnd && (max_steps < 0 || it < max_steps)) {
        CUDA_TRY(cudaMemsetAsync(d_max, 0, sizeof(unsigned long long), s->stream));
        absmax_kernel<REAL><<<grid, 256, 0, s->stream>>>((const REAL*)s->buf[s->cur], s->d.nx, (long long)s->pitch, rows, d_max);
        CUDA_TRY(cudaGetLastError());
        s->launches++;
        unsigned long long bits = 0;
        CUDA_TRY(cudaMemcpyAsync(&bits, d_max, sizeof(bits), cudaMemcpyDeviceToHost, s->stream));
        CUDA_TRY(cudaStreamSynchronize(s->stream));
        double umax_d;
        memcpy(&umax_d, &bits, sizeof(umax_d));
        const REAL umax = (REAL)umax_d;                     // exact: the reduction widened REAL values
        dt = CFL * dx / umax;
        if ((t + dt) > tEnd) { dt = tEnd - t; }
        t += dt; it += 1;
        int rc = do_step(s, (double)dt);
        if (rc) return rc;
    }
    if (dt_last) *dt_last = dt;
    if (steps) *steps = it;
    if (t_reached) *t_reached = t;
    return ADV_OK;
}

// ---- the same with everything on the device (Burgers 3-D through the pair kernel; any number of ranks) ----
// Stage 3 of every step folds max|u| of what it stores into the state block (no extra pass over the field); ranks
// combine their maxima with one 8-byte ncclAllReduce(max); a one-thread kernel turns the global maximum into dt,
// advances t and `it`, raises `done` once t has reached tEnd, and writes a status line the host can see.  The
// stage kernels read dt and `done` from the state block.  The host never waits for the step it has just enqueued:
// it enqueues step k as soon as it has SEEN the status of step k-2 (not done), so the GPU always has the next step
// queued, and exactly one step beyond the last real one gets enqueued -- on every rank alike, which keeps the halo
// flag counters of the ranks in step.  That last step's kernels see `done` and write nothing.
template <typename REAL>
__global__ void cfl_dt_kernel(CflDevState* st, CflHostStatus* hs, const unsigned long long* gmax)
{
    if (!st->done) {
        REAL t = (REAL)st->t, it = (REAL)st->it;
        const REAL tEnd = (REAL)st->tEnd;
        if (!(t < tEnd) || (st->max_steps >= 0 && !(it < st->max_steps))) {
            st->done = 1;
        } else {
            const REAL umax = (REAL)__longlong_as_double((long long)*gmax);
            REAL dt = (REAL)st->cfl * (REAL)st->dx / umax;          // LFWENO5FDM3d.m:71
            if ((t + dt) > tEnd) dt = tEnd - t;
            t += dt; it += 1;
            st->t = t; st->it = it; st->dt = dt;
            st->local_max = 0;                                      // stage 3 of this step starts a new maximum
        }
    }
    st->seq += 1;
    hs->t = st->t; hs->dt = st->dt; hs->it = st->it; hs->done = st->done;
    __threadfence_system();
    hs->seq = st->seq;
    __threadfence_system();
}

static bool burgers3d_pair_usable(const adv_solver* s)
{
    return (s->d.target == ADV_TARGET_BURGERS3D_MGPU || s->d.target == ADV_TARGET_BURGERS3D_SGPU) &&
           !env_on("ADV_BURGERS_DIRECT") && !env_on("ADV_BURGERS_RING") && get_tmap_encode() != nullptr && s->d.nx >= 8;
}

template <typename REAL>
static int run_until_cfl_device_impl(adv_solver* s, double tEnd_d, int max_steps, double* dt_last, double* steps, double* t_reached)
{
    if (!s->cfl_host) {
        CUDA_TRY(cudaHostAlloc((void**)&s->cfl_host, sizeof(CflHostStatus), cudaHostAllocMapped));
        CUDA_TRY(cudaHostGetDevicePointer((void**)&s->cfl_host_dev, (void*)s->cfl_host, 0));
    }
    memset((void*)s->cfl_host, 0, sizeof(CflHostStatus));
    CflDevState* cs = (CflDevState*)(s->arena + s->off_flags + 512);
    CflDevState h0;
    memset(&h0, 0, sizeof(h0));
    h0.tEnd = (double)(REAL)tEnd_d; h0.cfl = (double)(REAL)s->d.coef; h0.dx = (double)(REAL)s->dx; h0.max_steps = max_steps;
    CUDA_TRY(cudaMemcpyAsync(cs, &h0, sizeof(h0), cudaMemcpyHostToDevice, s->stream));
    // the first step's maximum: one reduction pass over the slab (later ones come out of stage 3)
    const long long rows = (long long)s->urows * s->nzs;
    const int grid = (int)(rows < current_sm_count() * 8 ? rows : current_sm_count() * 8);
    absmax_kernel<REAL><<<grid, 256, 0, s->stream>>>((const REAL*)s->buf[s->cur], s->d.nx, (long long)s->pitch, rows, &cs->local_max);
    CUDA_TRY(cudaGetLastError());
    s->launches++;
    const int cur0 = s->cur;
    s->cfl_on = true;
    int rc = ADV_OK;
    unsigned enq = 0;
    auto enqueue_step = [&]() -> int {
        const unsigned long long* gmax = &cs->local_max;
        if (s->d.nranks > 1) {
            int r = s->comm.allreduce_max_u64(&cs->local_max, &cs->global_max, s->stream, g_err, sizeof(g_err));
            if (r) return r;
            gmax = &cs->global_max;
        }
        cfl_dt_kernel<REAL><<<1, 1, 0, s->stream>>>(cs, s->cfl_host_dev, gmax);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return fail(ADV_ERR_CUDA, "cfl_dt_kernel: %s", cudaGetErrorString(e));
        s->launches++;
        enq++;
        return do_step(s, 0.0);         // dt comes from the state block
    };
    // steps 1 and 2 unconditionally; step k (k >= 3) once the status of step k-2 has been seen and is not `done`
    for (int k = 0; k < 2 && !rc; k++) rc = enqueue_step();
    while (!rc) {
        const unsigned need = enq - 1;                      // about to enqueue step enq+1: wants the status of step enq-1
        unsigned long long spins = 0;
        while (s->cfl_host->seq < need) {
            if ((++spins & 0xfffff) == 0 && cudaStreamQuery(s->stream) != cudaErrorNotReady && s->cfl_host->seq < need) {
                rc = fail(ADV_ERR_CUDA, "dynamic CFL run: the stream stopped without a status update (%s)", cudaGetErrorString(cudaGetLastError()));
                break;
            }
        }
        if (rc) break;
        // The status words may already belong to a later step than `need`.  `done` is sticky and a step is real
        // exactly when it raised `it`, so "step `need` was not real" <=> done && it < need -- whichever later
        // status is read, every rank derives the same answer and enqueues the same number of steps.
        if (s->cfl_host->done && (double)need > s->cfl_host->it) break;
        rc = enqueue_step();
    }
    s->cfl_on = false;
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    const double it = s->cfl_host->it;
    s->cur = (cur0 + (int)((long long)it % 3)) % 3;         // steps past the end wrote nothing
    if (dt_last) *dt_last = s->cfl_host->dt;
    if (steps) *steps = it;
    if (t_reached) *t_reached = s->cfl_host->t;
    return ADV_OK;
}

extern "C" int adv_run_until_cfl(adv_solver* s, double tEnd, int max_steps, double* dt_last, double* steps, double* t_reached)
{
    if (!s) return fail(ADV_ERR_INVALID, "NULL solver");
    if (!s->uploaded) return fail(ADV_ERR_STATE, "adv_run_until_cfl before adv_upload");
    if (is_heat(s->d.target)) return fail(ADV_ERR_INVALID, "the dynamic CFL step applies to the Burgers targets");
    const bool on_device = burgers3d_pair_usable(s) && !env_on("ADV_CFL_HOST_DT");
    if (s->d.nranks > 1 && !on_device)
        return fail(ADV_ERR_INVALID, "adv_run_until_cfl with nranks > 1 needs the 3-D Burgers pair kernel (device-resident time step)");
    CUDA_TRY(cudaSetDevice(s->d.device));
    CUDA_TRY(cudaEventRecord(s->ev_start, s->stream));
    int rc;
    if (on_device)
        rc = s->esz == 8 ? run_until_cfl_device_impl<double>(s, tEnd, max_steps, dt_last, steps, t_reached)
                         : run_until_cfl_device_impl<float>(s, tEnd, max_steps, dt_last, steps, t_reached);
    else
        rc = s->esz == 8 ? run_until_cfl_impl<double>(s, tEnd, max_steps, dt_last, steps, t_reached)
                         : run_until_cfl_impl<float>(s, tEnd, max_steps, dt_last, steps, t_reached);
    if (rc) return rc;
    if ((rc = halo_quiesce(s))) return rc;
    CUDA_TRY(cudaEventRecord(s->ev_stop, s->stream));
    return ADV_OK;
}

extern "C" int adv_solve_host(const adv_desc* desc, const void* host_in, void* host_out,
                              double tEnd_or_iters, double* dt_last, double* steps, double* t_reached)
{
    if (!desc || !host_in || !host_out) return fail(ADV_ERR_INVALID, "NULL argument");
    adv_desc d = *desc;
    d.nranks = 1; d.rank = 0;
    adv_solver* s = nullptr;
    int rc = adv_create(&d, &s);
    if (rc) return rc;
    rc = adv_upload(s, host_in);
    const bool iters_mode = is_heat(d.target) && !is_mgpu(d.target);
    if (!rc) {
        if (iters_mode) {
            rc = adv_step(s, (int)tEnd_or_iters);
            if (dt_last) *dt_last = s->dt_fixed;
            if (steps) *steps = (int)tEnd_or_iters;
            if (t_reached) *t_reached = s->dt_fixed * (int)tEnd_or_iters;
        } else {
            rc = adv_run_until(s, tEnd_or_iters, -1, dt_last, steps, t_reached);
        }
    }
    if (!rc) {
        if (is_mgpu(d.target)) {
            // Merge_domains (Tools.c:205-224): only owned planes are copied back; the six pad planes
            // of the output keep the values of the input.
            const size_t n = adv_local_elems(s), xy = (size_t)d.nx * s->urows;
            void* tmp = malloc(n * s->esz);
            if (!tmp) rc = fail(ADV_ERR_NOMEM, "out of host memory");
            if (!rc) rc = adv_download(s, tmp);
            if (!rc) {
                if (host_out != host_in) memcpy(host_out, host_in, n * s->esz);
                memcpy((char*)host_out + 3 * xy * s->esz, (char*)tmp + 3 * xy * s->esz, xy * (size_t)s->nzl * s->esz);
            }
            free(tmp);
        } else {
            rc = adv_download(s, host_out);
        }
    }
    char keep[sizeof(g_err)];
    memcpy(keep, g_err, sizeof(keep));
    adv_destroy(s);
    memcpy(g_err, keep, sizeof(keep));
    return rc;
}

extern "C" int adv_comm_unique_id(unsigned char id[128])
{
    if (!id) return fail(ADV_ERR_INVALID, "id is NULL");
    return HaloComm::unique_id(id, g_err, sizeof(g_err));
}

// ------------------------------------------------------------------------------------------
// Burgers stage without any allocation (adv_burgers3d_stage_f64_noalloc)
//
// Same kernel and same arithmetic as adv_burgers3d_stage_f64, but the j=2 row accumulator lives
// in device memory that is part of this library's image instead of memory somebody allocated.
// For hosts that may not allocate on the device at all: an embedding application that has
// handed the whole device budget to its own pool, a time-critical loop that cannot afford the
// device-wide synchronisation cudaMalloc/cudaFree imply, or a caller that simply owns no
// allocator.  Nothing here calls cudaMalloc, cudaFree or any library routine that allocates.
// ------------------------------------------------------------------------------------------
static __device__ __align__(256) double g_noalloc_rowacc[ADV_NOALLOC_MAX_ELEMS];

// In host code `g_noalloc_rowacc` is only a handle for the runtime, not an address the GPU can
// use; the device address has to be asked for, once per device, with that device current.
// The per-device slots are atomics: the C drivers run one host thread per GPU, and two threads may
// make their first call at the same moment (both then look the address up and store the same value).
// The cached addresses belong to the device's primary context: they do not survive cudaDeviceReset.
static double* noalloc_rowacc_for_device(int dev, int* rc_out)
{
    enum { MAX_DEV = 64 };
    static std::atomic<double*> ptr[MAX_DEV];       // zero-initialised; one slot per device ordinal
    *rc_out = ADV_OK;
    if (dev < 0 || dev >= MAX_DEV) { *rc_out = fail(ADV_ERR_INVALID, "device ordinal %d not supported", dev); return nullptr; }
    double* p = ptr[dev].load(std::memory_order_acquire);
    if (!p) {
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_rowacc);
        if (e != cudaSuccess || !p) {
            *rc_out = fail(ADV_ERR_CUDA, "cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e));
            return nullptr;
        }
        ptr[dev].store(p, std::memory_order_release);
    }
    return p;
}

extern "C" int adv_burgers3d_stage_f64_noalloc(const adv_stage_args* a, const double* u_in, const double* uo,
                                               double* u_out, void* cuda_stream)
{
    if (!a) return fail(ADV_ERR_INVALID, "NULL argument");
    if (a->target != ADV_TARGET_BURGERS3D_MGPU)
        return fail(ADV_ERR_INVALID, "adv_burgers3d_stage_f64_noalloc serves ADV_TARGET_BURGERS3D_MGPU only (the other "
                                     "targets use no scratch: call adv_burgers3d_stage_f64 with rowacc = NULL)");
    if (a->nzs <= 0 || a->pitch == 0) return fail(ADV_ERR_INVALID, "empty field");
    const unsigned long long need = (unsigned long long)a->pitch * (unsigned long long)a->nzs;
    if (need > (unsigned long long)ADV_NOALLOC_MAX_ELEMS)
        return fail(ADV_ERR_INVALID, "field needs %llu scratch elements (pitch %zu x %d planes) but the built-in scratch holds "
                                     "%llu; use adv_burgers3d_stage_f64 with caller-provided scratch",
                    need, a->pitch, a->nzs, (unsigned long long)ADV_NOALLOC_MAX_ELEMS);
    int rc, dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    cudaStream_t st = (cudaStream_t)cuda_stream;
    if (st != nullptr && st != cudaStreamLegacy && st != cudaStreamPerThread) {
        // the array whose address we hand to the kernel is the CURRENT device's copy: a stream of another
        // device would run the kernel there with a pointer into this one's memory
        int sdev = -1;
        cudaError_t e = cudaStreamGetDevice(st, &sdev);
        if (e != cudaSuccess) return fail(ADV_ERR_CUDA, "cudaStreamGetDevice: %s", cudaGetErrorString(e));
        if (sdev != dev)
            return fail(ADV_ERR_INVALID, "the stream belongs to device %d but device %d is current: call cudaSetDevice(%d) "
                                         "before every call of adv_burgers3d_stage_f64_noalloc", sdev, dev, sdev);
    }
    double* scratch = noalloc_rowacc_for_device(dev, &rc);
    if (!scratch) return rc;
    if (a->stage == 1 && a->k_begin >= 0 && a->k_end <= a->nzs && a->k_begin < a->k_end) {
        // the accumulator of a new time step starts from zero (the reference zeroes Lu once per step,
        // main.c:202); whatever an earlier call left in the shared array must not leak into this one.
        // Only the rows of the planes this call produces are cleared, so a stage split into several
        // plane windows does not wipe what an earlier window of the same stage wrote.
        CUDA_TRY(cudaMemsetAsync(scratch + a->pitch * (size_t)a->k_begin, 0,
                                 a->pitch * (size_t)(a->k_end - a->k_begin) * sizeof(double), st));
    }
    return burgers3d_stage_impl<double>(a, u_in, uo, u_out, scratch, st);
}

// ------------------------------------------------------------------------------------------
// peer (NVLink) halo exchange set-up
// ------------------------------------------------------------------------------------------
static int peer_supported(adv_solver* s)
{
    if (s->d.nranks < 2) return fail(ADV_ERR_INVALID, "peer halo exchange needs nranks > 1");
    if (s->d.target == ADV_TARGET_HEAT3D_MGPU) {
        const size_t vec = 16 / s->esz;
        if (s->d.nx % 2 || s->pitch % vec || !get_tmap_encode())
            return fail(ADV_ERR_INVALID, "fused halo stores need the TMA heat kernel (even nx)");
    }
    return ADV_OK;
}

extern "C" int adv_peer_handle(adv_solver* s, unsigned char handle[64])
{
    if (!s || !handle) return fail(ADV_ERR_INVALID, "NULL argument");
    CUDA_TRY(cudaSetDevice(s->d.device));
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaIpcGetMemHandle(&h, s->arena));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle, &h, 64);
    return ADV_OK;
}

extern "C" int adv_peer_connect(adv_solver* s, const unsigned char* left_handle, const unsigned char* right_handle)
{
    if (!s) return fail(ADV_ERR_INVALID, "NULL solver");
    int rc = peer_supported(s);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(s->d.device));
    const unsigned char* hs[2] = {left_handle, right_handle};
    const bool need[2] = {s->d.rank > 0, s->d.rank < s->d.nranks - 1};
    for (int n = 0; n < 2; n++) {
        if (!need[n]) continue;
        if (!hs[n]) return fail(ADV_ERR_INVALID, "rank %d needs its %s neighbour's handle", s->d.rank, n ? "right" : "left");
        cudaIpcMemHandle_t h;
        memcpy(&h, hs[n], 64);
        void* p = nullptr;
        CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        s->peer_base[n] = (char*)p;
        s->peer_is_ipc[n] = true;
    }
    s->p2p = 1;
    return ADV_OK;
}

extern "C" int adv_peer_connect_local(adv_solver* s, adv_solver* left, adv_solver* right)
{
    if (!s) return fail(ADV_ERR_INVALID, "NULL solver");
    int rc = peer_supported(s);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(s->d.device));
    adv_solver* nb[2] = {left, right};
    const bool need[2] = {s->d.rank > 0, s->d.rank < s->d.nranks - 1};
    for (int n = 0; n < 2; n++) {
        if (!need[n]) continue;
        if (!nb[n]) return fail(ADV_ERR_INVALID, "rank %d needs its %s neighbour", s->d.rank, n ? "right" : "left");
        if (nb[n]->arena_bytes != s->arena_bytes) return fail(ADV_ERR_INVALID, "neighbour slab has a different layout");
        int can = 0;
        CUDA_TRY(cudaDeviceCanAccessPeer(&can, s->d.device, nb[n]->d.device));
        if (!can) return fail(ADV_ERR_CUDA, "device %d cannot access device %d", s->d.device, nb[n]->d.device);
        cudaError_t e = cudaDeviceEnablePeerAccess(nb[n]->d.device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(ADV_ERR_CUDA, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        s->peer_base[n] = nb[n]->arena;
        s->peer_is_ipc[n] = false;
    }
    s->p2p = 1;
    return ADV_OK;
}

extern "C" int adv_halo_mode(const adv_solver* s)
{
    if (!s || s->d.nranks < 2) return 0;
    return s->p2p ? 2 : 1;
}
