// burgers3d_stage_pair: the fused 3-D Burgers stage (LF split + WENO5 JS/Z faces in x, y, z + divergence
// [+ viscous term] + SSP-RK3 combine), third generation.  Same formulas as burgers3d_stage_ring (pair formulation,
// (f+,f-) plane ring in shared memory) with a leaner side routine (weno_core6, burgers_kernels.cuh: 40 FP64
// instructions instead of weno_side6's 43); what changed most is everything around the arithmetic, because the kernel
// is bound by the issue port (an FP64 instruction holds it two cycles, anything else one: DESIGN.md 3.2) and 31 % of
// the ring kernel's steady-state instructions were not FP64:
//
//  * TMA staging.  Per plane ONE cp.async.bulk.tensor.3d box brings u of tile + halo (40 x 22 values, x halo
//    4 so that the box origin falls on 16 bytes) into shared memory, issued by one elected thread a whole
//    iteration ahead and completing on an mbarrier; points outside the array (and planes outside the slab)
//    arrive as zeros, so no thread carries halo offsets, bounds predicates or 64-bit address arithmetic for
//    the staging, in interior and rim tiles alike.  Threads then split the raw values once into (f+, f-) pairs
//    ({3/4,-1/4} u^2 by sign, exact) in the ring slot.  The chunk's five run-in planes are five boxes in
//    flight at once on one barrier.
//  * y-adjacent cell pairs.  A thread owns cells (tx, 2ty) and (tx, 2ty+1): the two y stencils share four of
//    their six rows (4 neighbour loads instead of 8), the face between the two cells needs no exchange, and
//    only one half per cell crosses to the neighbouring thread through shared memory (2 stores + 2 loads per
//    thread and plane instead of 4 + 8).
//  * Centred z pairs.  (hn_z, hp_z) of plane p come from the thread's own ring column p-2..p+2 (which the
//    thread staged itself: no barrier in front of it); the face flux is hn_z(p-1) [carried] + hp_z(p).
//  * No extra warp.  The 96 halves on the tile rim (64 in y, 32 in x) are three warp-rounds per plane, taken by
//    three of the eight compute warps in rotation ((warp + plane) mod 8), so no scheduler carries a ninth warp
//    and all four see the same load over time.
//  * Steady state: planes for which every window / ownership / staging test is true run a body compiled without
//    them (STEADY), unrolled by two so that the buffers that alternate from plane to plane have compile-time indices;
//    the x-rim halves are computed one plane ahead, so the finalisation has no special lanes.  The finalisation of rim
//    tiles (cells next to the domain faces, the i = 2 / j = 2 pseudo-updates, the j = 2 scratch row) is compiled
//    separately (RIM), chosen per CTA.
//  * Multi-rank slabs: the launch's first z chunks can be the boundary windows of a fused halo exchange (SlabSched,
//    common.cuh); a dynamic-CFL run reads dt / done from device memory and (TRACK) folds max|u| into stage 3.
// Reference: Compute_dF/dG/dH + Reconstruct1d + Compute_RK, MultiGPU/Burgers3d_Baseline/Kernels.cu:112-486;
// WENO5Zreconstruction + Compute_Laplace_Async + Compute_RK, SingleGPU/Burgers3d_WENO5/kernels.cu:142-530.
#pragma once
#include "burgers_kernels.cuh"
#include "heat_kernels.cuh"     // tma_load_3d, mbar_expect_tx, mbar_wait_parity
#include <type_traits>

namespace adv {

template <typename REAL>
struct BurgersPairSmem {
    static constexpr int TX = 32, TY = 16, H = 3, HXR = 4, NW = TY / 2;
    static constexpr int RP = TX + 2 * HXR, ROWS = TY + 2 * H, NPTS = RP * ROWS;       // 40 x 22 = 880 points per plane
    static constexpr int NSLOT = 6;                                                     // planes p-2 .. p+2 live, p+3 being staged
    static constexpr int BOXB = NPTS * (int)sizeof(REAL);                               // bytes one TMA box delivers
    static constexpr int RAWB = (BOXB + 127) / 128 * 128;
    typedef typename V2T<REAL>::type V2;
    struct alignas(128) Slot { V2 v[NPTS]; };
    struct alignas(128) Raw { unsigned char b[RAWB]; };
    // halves that cross between threads of adjacent row pairs: hnU[.][t+1] = hn_y of the UPPER cell of row pair t
    // (index 0: the row below the tile), hpL[.][t] = hp_y of the LOWER cell of row pair t (index NW: the row above)
    struct alignas(128) Exch { REAL hnU[2][NW + 1][TX], hpL[2][NW + 1][TX]; };
    Slot slot[NSLOT];
    Raw raw[2];
    Exch ex;
    alignas(16) REAL edge_hn_x[2][TY], edge_hp_x[2][TY];    // hn of column -1, hp of column TX (rows 2t, 2t+1 read as one pair)
    unsigned long long bar, full[2], pro;
    REAL dt;                                    // the time step: the launch argument, or the device-resident one of a dynamic CFL run
    static_assert(sizeof(Exch) >= RAWB, "the exchange block doubles as a landing zone of the run-in planes");
    // landing zones of the chunk's five run-in planes: raw[0], raw[1], ex, and slot[5] (one or two boxes)
    static constexpr int NLAND = (2 * RAWB <= (int)sizeof(Slot)) ? 5 : 4;
};

// TRACK: fold max|u_out| into *a.cfl_umax (stage 3 of a dynamic CFL run); a separate instantiation so that the
// fixed-dt stage 3 carries none of it
template <typename REAL, int W, int MODE, int STAGE, bool PEER, bool TRACK = false>
__global__ void __launch_bounds__(256, 2)
burgers3d_stage_pair(const __grid_constant__ CUtensorMap tmap, BurgersArgs<REAL> a, SlabSched sched,
                     const REAL* __restrict__ u, const REAL* __restrict__ uo, REAL* __restrict__ out, REAL* __restrict__ rowacc)
{
    typedef BurgersPairSmem<REAL> S;
    typedef typename S::V2 V2;
    constexpr int TX = S::TX, TY = S::TY, H = S::H, HXR = S::HXR, RP = S::RP, NSLOT = S::NSLOT, NW = S::NW, NLAND = S::NLAND;
    extern __shared__ __align__(128) unsigned char pair_smem_raw[];
    S& sm = *reinterpret_cast<S*>(pair_smem_raw);

    const int tx = threadIdx.x, ty = threadIdx.y;
    const int tid = ty * 32 + tx;
    const int i0 = blockIdx.x * TX, j0 = blockIdx.y * TY;
    int kb, ke;
    const int bw = slab_window(sched, kb, ke);          // boundary window (fused halo exchange) or interior chunk
    if (kb >= ke) return;
    REAL* const out2 = (REAL*)slab_out2(sched, bw, a.out2);    // PEER: second destination of every store (null: none)
    const long long off2 = slab_off2(sched, bw, a.off2);
    if (a.cfl_done && *a.cfl_done) {        // dynamic CFL run that has reached tEnd: a speculatively enqueued step writes nothing,
        slab_signal(sched, threadIdx.x == 0 && threadIdx.y == 0);   // but keeps the neighbours' stage counters moving
        return;
    }
    constexpr bool track = TRACK;
    unsigned long long umax_bits = 0;       // max |value stored| by this thread (stage 3 of a dynamic CFL run)
    const long long P = a.pitch, Z = a.plane;
    const REAL* __restrict__ wc = a.wc;
    const unsigned bar = (unsigned)__cvta_generic_to_shared(&sm.bar);
    const unsigned pro = (unsigned)__cvta_generic_to_shared(&sm.pro);
    const unsigned full0 = (unsigned)__cvta_generic_to_shared(&sm.full[0]);
    const bool interior = (i0 >= 3) && (i0 + TX - 1 <= a.nx - 4) && (j0 >= 3) && (j0 + TY - 1 <= a.ny - 4);

    if (tid == 0) {
        sm.dt = a.cfl_done ? (REAL)*a.cfl_dt : a.dt;    // read from shared memory where it is used: no register held across the loop
        mbar_init(bar, NW);              // one arrival per warp: lane 0, after __syncwarp()
        mbar_init(pro, 1);
        mbar_init(full0, 1);
        mbar_init(full0 + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    slab_wait(sched, bw, tid == 0);                     // the neighbour's previous stage has landed in our ghost planes

    // ---- what this thread stages per plane: its two cells and one or two points of the frame around the tile
    const int ownL = (2 * ty + H) * RP + tx + HXR, ownU = ownL + RP;
    int hidx[2];
#pragma unroll
    for (int n = 0; n < 2; n++) {
        const int hh = tid + n * 256;
        int r, c;
        if (hh < H * RP) { r = hh / RP; c = hh - r * RP; }
        else if (hh < 2 * H * RP) { const int t = hh - H * RP; r = t / RP; c = t - r * RP; r += H + TY; }
        else { const int t = hh - 2 * H * RP; r = H + t / (2 * HXR); const int q = t - (t / (2 * HXR)) * (2 * HXR); c = q < HXR ? q : TX + q; }
        hidx[n] = (hh < S::NPTS - TX * TY) ? r * RP + c : -1;
    }
    auto split = [&](REAL v) -> V2 {
        const REAL sq = v * v;
        const bool ng = is_neg<REAL>(v);
        V2 r;
        r.x = (ng ? REAL(-0.25) : REAL(0.75)) * sq;     // f+ = 0.5*(u^2/2 + |u|u)
        r.y = (ng ? REAL(0.75) : REAL(-0.25)) * sq;     // f- = 0.5*(u^2/2 - |u|u)
        return r;
    };
    auto split_plane = [&](const void* rawp, int sidx) {
        const REAL* __restrict__ rp = reinterpret_cast<const REAL*>(rawp);
        V2* __restrict__ sp = sm.slot[sidx].v;
        const REAL vL = rp[ownL], vU = rp[ownU], v0 = rp[hidx[0]];
        REAL v1 = REAL(0);
        if (hidx[1] >= 0) v1 = rp[hidx[1]];
        sp[ownL] = split(vL);
        sp[ownU] = split(vU);
        sp[hidx[0]] = split(v0);
        if (hidx[1] >= 0) sp[hidx[1]] = split(v1);
    };
    auto slot_of = [&](int s0, int d) -> int { int t = s0 + d; return t >= NSLOT ? t - NSLOT : t; };
    const REAL* __restrict__ we = (W == WENO_JS) ? a.wp : a.wc;
    // the two halves centred on c along a line (m2, m1, c, p1, p2): hn from the f+ components, hp from the f- ones
    auto half_n = [&](REAL m2, REAL m1, REAL c, REAL p1, REAL p2) -> REAL {
        const REAL g0 = m1 - m2, g1 = c - m1, g2 = p1 - c, g3 = p2 - p1;
        return weno_core6<REAL, W, false>(c, g0, g1, g2, g3, g1 - g0, g2 - g1, g3 - g2, we, wc);
    };
    auto half_p = [&](REAL m2, REAL m1, REAL c, REAL p1, REAL p2) -> REAL {
        const REAL g0 = m1 - m2, g1 = c - m1, g2 = p1 - c, g3 = p2 - p1;
        return weno_core6<REAL, W, true>(c, g0, g1, g2, g3, g1 - g0, g2 - g1, g3 - g2, we, wc);
    };

    // ---- run-in: planes kb-3 .. kb+1 into slots 0 .. 4, all boxes of a round in flight on one barrier
    {
        unsigned char* zone[5] = {sm.raw[0].b, sm.raw[1].b, reinterpret_cast<unsigned char*>(&sm.ex),
                                  reinterpret_cast<unsigned char*>(&sm.slot[5]), reinterpret_cast<unsigned char*>(&sm.slot[5]) + S::RAWB};
        unsigned ppar = 0;
        for (int d0 = 0; d0 < 5; d0 += NLAND) {
            const int nd = min(NLAND, 5 - d0);
            if (tid == 0) {
                mbar_expect_tx(pro, (unsigned)(nd * S::BOXB));
                for (int d = 0; d < nd; d++)
                    tma_load_3d((unsigned)__cvta_generic_to_shared(zone[d]), &tmap, i0 - HXR, j0 - H, kb - 3 + d0 + d, pro);
            }
            mbar_wait_parity(pro, ppar);
            ppar ^= 1u;
            for (int d = 0; d < nd; d++) split_plane(zone[d], d0 + d);
            __syncthreads();
        }
    }
    if (tid == 0) {     // plane kb+2: split by the first iteration (n = 0) out of raw[1]
        mbar_expect_tx(full0 + 8, (unsigned)S::BOXB);
        tma_load_3d((unsigned)__cvta_generic_to_shared(sm.raw[1].b), &tmap, i0 - HXR, j0 - H, kb + 2, full0 + 8);
    }

    // ---- per-cell bookkeeping of rim tiles: which directional contributions a cell takes
    const int i = i0 + tx;
    const int jL = j0 + 2 * ty;
    bool in_dom[2];
    unsigned act = 0;
    {
        const bool i_in3 = (i >= 3 && i <= a.nx - 4);
#pragma unroll
        for (int c = 0; c < 2; c++) {
            const int j = jL + c;
            in_dom[c] = i < a.nx && j < a.ny;
            const bool j_in3 = (j >= 3 && j <= a.ny - 4);
            const bool xa = (MODE == BMODE_MGPU) ? (j_in3 && i >= 2 && i <= a.nx - 4) : (i_in3 && j_in3);   // Kernels.cu:250,266
            const bool ya = (MODE == BMODE_MGPU) ? (i_in3 && j >= 2 && j <= a.ny - 4) : (i_in3 && j_in3);   // Kernels.cu:323,339
            const bool za = i_in3 && j_in3;                                                                 // Kernels.cu:397
            const bool row2 = (MODE == BMODE_MGPU) && j == 2 && ya;
            act |= ((xa ? 1u : 0u) | (ya ? 2u : 0u) | (za ? 4u : 0u) | (row2 ? 8u : 0u)) << (4 * c);
        }
    }
    const bool x_seed0 = (MODE == BMODE_MGPU) && i == 2;
    long long col0 = (long long)i + P * jL;             // cell c lives at col0 + c*P
    if (!in_dom[0]) col0 = 0;

    auto visc13 = [&](const REAL* __restrict__ pc, REAL c0) -> REAL {
        return a.kx * (-pc[-2] + 16 * pc[-1] - 30 * c0 + 16 * pc[1] - pc[2]) +
               a.ky * (-pc[-2 * P] + 16 * pc[-P] - 30 * c0 + 16 * pc[P] - pc[2 * P]) +
               a.kz * (-pc[-2 * Z] + 16 * pc[-Z] - 30 * c0 + 16 * pc[Z] - pc[2 * Z]);
    };

    // planes whose z halves are needed: every finalised owned plane q needs the faces q-1/2 and q+1/2
    const int zlo = max(kb, a.own_lo) - 1, zhi = min(ke, a.own_hi);

    // carried from the iteration that computed the pairs of plane p-1
    REAL X1[2] = {REAL(0), REAL(0)}, X2[2] = {REAL(0), REAL(0)};    // 6 x the x face fluxes right / left of the cell
    REAL hpLow = REAL(0), Yint = REAL(0), hnUp = REAL(0);           // own y halves: hp_y(L), hn_y(L)+hp_y(U), hn_y(U)
    REAL hn_prev[2] = {REAL(0), REAL(0)}, Fz_prev[2] = {REAL(0), REAL(0)};

    // plane barrier: a warp arrives once (lane 0, after __syncwarp() has ordered the other lanes' shared-memory
    // writes before it) -- 8 arrivals per phase instead of 256 wake the waiting warps far less often
    auto warp_arrive = [&]() { __syncwarp(); if (tx == 0) mbar_arrive(bar); };
    warp_arrive();
    unsigned fpar = 0;
    int sc = 2;                                          // ring slot of plane p (slot of plane q = (q - (kb-3)) mod 6)
    // running element offset of cell L in the plane being finished (plane p-1), bumped once per iteration
    long long om = col0 + Z * (kb - 3);

    // One iteration: centre plane p = kb-1+n.  Buffers that alternate from plane to plane are indexed by the
    // parity of n (NPAR: compile-time 0/1 in the unrolled steady-state loop, 2 = take it from n).
    //   ex[n&1], edge_x[n&1]      halves of plane p that cross threads (written in iteration n / n-1, read in n+1 / n)
    //   raw[(n+1)&1]              box of plane p+3 (issued in iteration n-1, split in iteration n)
    // STEADY: every plane-level test is known to be true (plane p-1 is finished and owned, planes p and p+1 are
    // owned and inside the chunk, planes p+3 and p+4 are still to be staged) and is compiled out.
    auto iteration = [&](auto rim_tag, auto steady_tag, auto par_tag, int p, int n) {
        constexpr bool RIM = decltype(rim_tag)::value;
        constexpr bool STEADY = decltype(steady_tag)::value;
        constexpr int NPAR = decltype(par_tag)::value;
        const int nb = (NPAR == 2) ? (n & 1) : NPAR, eb = nb ^ 1;
        const int m = p - 1;                                                    // the plane this iteration finishes
        const bool fin = STEADY || m >= kb;
        const bool own_m = STEADY || (fin && (m >= a.own_lo && m < a.own_hi));
        const bool do_z = STEADY || (p >= zlo && p <= zhi);
        const bool do_xy = STEADY || ((p >= kb) && (p < ke) && (p >= a.own_lo && p < a.own_hi));
        const bool do_xrim = STEADY || ((p + 1 < ke) && (p + 1 >= a.own_lo && p + 1 < a.own_hi));
        const bool do_stage = STEADY || (p + 3 <= ke + 2);
        const bool do_issue = STEADY || (p + 4 <= ke + 2);
        om += Z;

        // (A) u and uo of the plane being finished (L2 hits: the boxes brought them in three planes ago)
        REAL u_m[2] = {REAL(0), REAL(0)}, uo_m[2] = {REAL(0), REAL(0)};
        if (fin) {
#pragma unroll
            for (int c = 0; c < 2; c++)
                if (!RIM || in_dom[c]) {
                    u_m[c] = u[om + c * P];
                    if (STAGE > 1) uo_m[c] = uo[om + c * P];
                }
        }

        // (B) z halves of plane p from the thread's own ring column (staged by this thread: nothing to wait for)
        const V2* __restrict__ s = sm.slot[sc].v;
        const V2 cL = s[ownL], cU = s[ownU];
        REAL Fz[2] = {REAL(0), REAL(0)};
        if (do_z) {
            const V2 *q0 = sm.slot[slot_of(sc, 4)].v, *q1 = sm.slot[slot_of(sc, 5)].v,
                     *q3 = sm.slot[slot_of(sc, 1)].v, *q4 = sm.slot[slot_of(sc, 2)].v;
#pragma unroll
            for (int c = 0; c < 2; c++) {
                const int own = ownL + c * RP;
                const V2 v0 = q0[own], v1 = q1[own], v2 = c ? cU : cL, v3 = q3[own], v4 = q4[own];
                const REAL hn = half_n(v0.x, v1.x, v2.x, v3.x, v4.x);
                const REAL hp = half_p(v0.y, v1.y, v2.y, v3.y, v4.y);
                Fz[c] = hn_prev[c] + hp;                 // 6 x the face flux at p-1/2
                hn_prev[c] = hn;
            }
        }

        mbar_wait_hint(bar, (unsigned)nb);               // phase n: completed by the arrivals that end iteration n-1 (n = 0: by the arrivals before the loop)
        // everybody is done with the raw buffer of plane p+2: the box of plane p+4 may land there
        if (tid == 0 && do_issue) {
            const unsigned fb = full0 + 8 * nb;
            mbar_expect_tx(fb, (unsigned)S::BOXB);
            tma_load_3d((unsigned)__cvta_generic_to_shared(sm.raw[nb].b), &tmap, i0 - HXR, j0 - H, p + 4, fb);
        }

        // (E) finish plane m: divergence from the carried own halves + the neighbours' halves, RK combine, store
        if (fin) {
#pragma unroll
            for (int c = 0; c < 2; c++) {
                const int r = 2 * ty + c;
                const long long o = om + c * P;
                REAL Lu = REAL(0);
                REAL fu = REAL(0), fu_old = REAL(0), gu = REAL(0), gu_old = REAL(0);
                if (own_m) {
                    fu = X1[c];
                    fu_old = X2[c];
                    if (c == 0) { gu = Yint; gu_old = sm.ex.hnU[eb][ty][tx] + hpLow; }
                    else        { gu = hnUp + sm.ex.hpL[eb][ty + 1][tx]; gu_old = Yint; }
                }
                if (!RIM) {
                    if (own_m) {
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
                    const REAL res = rk_stage<REAL, STAGE>(u_m[c], uo_m[c], Lu, sm.dt, wc);
                    const REAL val = rk ? res : u_m[c];
                    out[o] = val;
                    if (PEER && out2) out2[o + off2] = val;
                    if (track) { const unsigned long long vb = (unsigned long long)__double_as_longlong(fabs((double)val)); umax_bits = vb > umax_bits ? vb : umax_bits; }
                } else {
                    const bool xa = (act >> (4 * c)) & 1u, ya = (act >> (4 * c)) & 2u, za = (act >> (4 * c)) & 4u;
                    const bool row2 = (act >> (4 * c)) & 8u;
                    if (own_m) {
                        fu_old = x_seed0 ? REAL(0) : fu_old;                                   // Kernels.cu:254
                        gu_old = ((MODE == BMODE_MGPU) && (j0 + r == 2)) ? REAL(0) : gu_old;   // Kernels.cu:327
                        Lu = xa ? -(fu - fu_old) * a.rdx6 : REAL(0);
                        Lu = ya ? Lu - (gu - gu_old) * a.rdy6 : Lu;
                    }
                    if (in_dom[c]) {
                        const bool rk = (MODE == BMODE_MGPU) ? true : (own_m && za);
                        if (own_m) {
                            if (row2) {     // the j = 2 row keeps accumulating within a time step (Kernels.cu:349, main.c:202)
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
                        const REAL res = rk_stage<REAL, STAGE>(u_m[c], uo_m[c], Lu, sm.dt, wc);
                        const REAL val = rk ? res : u_m[c];
                        out[o] = val;
                        if (PEER && out2) out2[o + off2] = val;
                        if (track) { const unsigned long long vb = (unsigned long long)__double_as_longlong(fabs((double)val)); umax_bits = vb > umax_bits ? vb : umax_bits; }
                    }
                }
            }
        }
#pragma unroll
        for (int c = 0; c < 2; c++) Fz_prev[c] = Fz[c];

        // (F) x and y halves of plane p, centred on the two own cells
        if (do_xy) {
            {
                const V2 ya_ = s[ownL - 2 * RP], yb = s[ownL - RP], ye = s[ownU + RP], yf = s[ownU + 2 * RP];
                // six rows, two cells: the five forward differences and four second differences per component are shared
                const REAL G0 = yb.x - ya_.x, G1 = cL.x - yb.x, G2 = cU.x - cL.x, G3 = ye.x - cU.x, G4 = yf.x - ye.x;
                const REAL T0 = G1 - G0, T1 = G2 - G1, T2 = G3 - G2, T3 = G4 - G3;
                const REAL hnL_ = weno_core6<REAL, W, false>(cL.x, G0, G1, G2, G3, T0, T1, T2, we, wc);
                const REAL hnU_ = weno_core6<REAL, W, false>(cU.x, G1, G2, G3, G4, T1, T2, T3, we, wc);
                const REAL H0 = yb.y - ya_.y, H1 = cL.y - yb.y, H2 = cU.y - cL.y, H3 = ye.y - cU.y, H4 = yf.y - ye.y;
                const REAL S0 = H1 - H0, S1 = H2 - H1, S2 = H3 - H2, S3 = H4 - H3;
                const REAL hpL_ = weno_core6<REAL, W, true>(cL.y, H0, H1, H2, H3, S0, S1, S2, we, wc);
                const REAL hpU_ = weno_core6<REAL, W, true>(cU.y, H1, H2, H3, H4, S1, S2, S3, we, wc);
                sm.ex.hnU[nb][ty + 1][tx] = hnU_;
                sm.ex.hpL[nb][ty][tx] = hpL_;
                hpLow = hpL_; Yint = hnL_ + hpU_; hnUp = hnU_;
            }
            // the halves beyond the tile's x rim were computed one iteration ago: lanes 0 and 31 take them in
            // place of the shuffled neighbour half, so the face sums are complete here and the finalisation
            // has no special lanes
            const V2 eN = *reinterpret_cast<const V2*>(&sm.edge_hn_x[nb][2 * ty]);
            const V2 eP = *reinterpret_cast<const V2*>(&sm.edge_hp_x[nb][2 * ty]);
#pragma unroll
            for (int c = 0; c < 2; c++) {
                const int own = ownL + c * RP;
                const V2 cc = c ? cU : cL;
                const V2 xm2 = s[own - 2], xm1 = s[own - 1], xp1 = s[own + 1], xp2 = s[own + 2];
                const REAL hn_x = half_n(xm2.x, xm1.x, cc.x, xp1.x, xp2.x);
                const REAL hp_x = half_p(xm2.y, xm1.y, cc.y, xp1.y, xp2.y);
                REAL hpR = __shfl_down_sync(0xffffffffu, hp_x, 1);
                REAL hnL = __shfl_up_sync(0xffffffffu, hn_x, 1);
                hpR = (tx == 31) ? (c ? eP.y : eP.x) : hpR;
                hnL = (tx == 0) ? (c ? eN.y : eN.x) : hnL;
                X1[c] = hn_x + hpR;
                X2[c] = hnL + hp_x;
            }
        }
        // (G) the halves on the tile rim: three warp-rounds, taken in rotation -- y rim of plane p, x rim of plane p+1
        {
            const int role = (ty + n) & 7;
            if ((role < 2 && do_xy) || (role == 2 && do_xrim)) {
                int oc, st;
                bool minus;
                if (role == 0)      { oc = (H - 1) * RP + tx + HXR;  st = RP; minus = true;  }              // hn_y of the row below
                else if (role == 1) { oc = (H + TY) * RP + tx + HXR; st = RP; minus = false; }              // hp_y of the row above
                else if (tx < TY)   { oc = (tx + H) * RP + HXR - 1;  st = 1;  minus = true;  }              // hn_x of column -1
                else                { oc = (tx - TY + H) * RP + HXR + TX; st = 1; minus = false; }          // hp_x of column TX
                const V2* __restrict__ sr = (role == 2) ? sm.slot[slot_of(sc, 1)].v : s;
                const V2 m2 = sr[oc - 2 * st], m1 = sr[oc - st], cc = sr[oc], p1 = sr[oc + st], p2 = sr[oc + 2 * st];
                const REAL v = half_n(minus ? m2.x : p2.y, minus ? m1.x : p1.y, minus ? cc.x : cc.y,
                                      minus ? p1.x : m1.y, minus ? p2.x : m2.y);
                if (role == 0) sm.ex.hnU[nb][0][tx] = v;
                else if (role == 1) sm.ex.hpL[nb][NW][tx] = v;
                else if (tx < TY) sm.edge_hn_x[eb][tx] = v;
                else sm.edge_hp_x[eb][tx - TY] = v;
            }
        }

        // (C) plane p+3: its box was issued an iteration ago
        if (do_stage) {
            mbar_wait_parity(full0 + 8 * eb, (fpar >> eb) & 1u);
            fpar ^= 1u << eb;
            split_plane(sm.raw[eb].b, slot_of(sc, 3));
        }
        warp_arrive();
        sc = (sc + 1 == NSLOT) ? 0 : sc + 1;
    };

    // steady-state centre planes: see STEADY above
    const int s_lo = max(kb, a.own_lo) + 1, s_hi = min(min(ke - 2, a.own_hi - 2), a.nzs - 5);
    auto run = [&](auto rim_tag) {
        typedef std::integral_constant<int, 0> P0;
        typedef std::integral_constant<int, 1> P1;
        typedef std::integral_constant<int, 2> PR;
        int p = kb - 1, n = 0;
        for (; p <= ke && (p < s_lo || (n & 1)); p++, n++) iteration(rim_tag, BoolTag<false>(), PR(), p, n);
        for (; p + 1 <= s_hi; p += 2, n += 2) {
            iteration(rim_tag, BoolTag<true>(), P0(), p, n);
            iteration(rim_tag, BoolTag<true>(), P1(), p + 1, n + 1);
        }
        for (; p <= ke; p++, n++) iteration(rim_tag, BoolTag<false>(), PR(), p, n);
    };
    if (interior) run(BoolTag<false>()); else run(BoolTag<true>());
    if (track) {                            // one atomic per warp
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, umax_bits, o);
            umax_bits = other > umax_bits ? other : umax_bits;
        }
        if (tx == 0) atomicMax(a.cfl_umax, umax_bits);
    }
    slab_signal(sched, tid == 0);                   // last CTA of a boundary window raises the neighbour's flag
}

}  // namespace adv
