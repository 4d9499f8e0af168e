/* Burgers3d.run / Diffusion3d.run (multi-GPU) -- drop-ins for MultiGPU/Burgers3d_Baseline/main.c and
 * MultiGPU/Diffusion3d_Baseline/main.c: same argv, stdout block, initial.bin and result.bin (float32,
 * z-padded to Nz+6 planes, pad planes holding the initial values).
 *   ADV_NGPUS=P ./Burgers3d.run  tEnd CFL L W H Nx Ny Nz block_x block_y block_z      (-DBURGERS)
 *   ADV_NGPUS=P ./Diffusion3d.run K L W H Nx Ny Nz max_iters block_x block_y block_z
 * `mpirun -np P` becomes ADV_NGPUS=P: ranks are GPUs of this node, one host thread per GPU, halo planes
 * travel GPU-to-GPU through NCCL.  Nz must be divisible by P (the reference silently drops planes). */
#include <pthread.h>
#include "common.h"

#define RADIUS ADV_RADIUS
typedef struct rank_ctx {
    adv_desc d;
    adv_solver* solver;
    struct rank_ctx* all;
    double tEnd;
    int iters;
    double* h_u;            /* global padded array (shared) */
    unsigned _Nz;
    double dt, it, t, HtD, DtH, compute;
    int rc;
    pthread_barrier_t* bar;
} rank_ctx;

static void* rank_main(void* p)
{
    rank_ctx* c = (rank_ctx*)p;
    const size_t xy = (size_t)c->d.nx * c->d.ny;
    adv_solver* s = NULL;
    c->rc = adv_create(&c->d, &s);
    if (c->rc) { printf("rank %d: %s\n", c->d.rank, adv_last_error()); die_now(); }
    printf("Process %d -> GPU%d\n", c->d.rank, c->d.device);
    /* all ranks live in this process: let the boundary kernels store into the neighbours' ghost planes
     * directly (NVLink peer access); ADV_HALO=nccl keeps the ncclSend/ncclRecv path */
    c->solver = s;
    pthread_barrier_wait(c->bar);
    if (c->d.nranks > 1 && !(getenv("ADV_HALO") && !strcmp(getenv("ADV_HALO"), "nccl"))) {
        adv_solver* left = c->d.rank > 0 ? c->all[c->d.rank - 1].solver : NULL;
        adv_solver* right = c->d.rank < c->d.nranks - 1 ? c->all[c->d.rank + 1].solver : NULL;
        if (adv_peer_connect_local(s, left, right)) printf("rank %d: peer halo unavailable (%s), using NCCL\n", c->d.rank, adv_last_error());
    }
    pthread_barrier_wait(c->bar);
    /* every rank must have ended up on the same exchange path: a mix would wait forever */
    for (int r = 0; r < c->d.nranks; r++)
        if (adv_halo_mode(c->all[r].solver) != adv_halo_mode(s)) { printf("ranks disagree on the halo exchange path\n"); die_now(); }
    const size_t n = adv_local_elems(s);
    double* h_s_u = (double*)malloc(sizeof(double) * n);

    die_on(adv_init_reference_ic(s), "adv_init_reference_ic");   /* Init_domain + Init_subdomain + H2D */
    die_on(adv_download(s, h_s_u), "adv_download");
    /* assemble the global initial condition (every rank stores planes [r*_Nz, r*_Nz+_Nz+6)) */
    memcpy(c->h_u + xy * (size_t)c->d.rank * c->_Nz, h_s_u, sizeof(double) * n);
    pthread_barrier_wait(c->bar);
    if (c->d.rank == 0) {
        save_binary_f32(c->h_u, xy * (c->d.nz + 2 * RADIUS), "initial.bin");
        printf("IC saved in Host rank %d\n", 0);
    }
    double t0 = now_s();
    die_on(adv_upload(s, h_s_u), "adv_upload");
    die_on(adv_sync(s), "adv_sync");
    c->HtD = now_s() - t0;

    pthread_barrier_wait(c->bar);
    t0 = now_s();
#ifdef BURGERS
    die_on(adv_run_until(s, c->tEnd, -1, &c->dt, &c->it, &c->t), "adv_run_until");
#else
    die_on(adv_run_until(s, adv_dt(s) * c->iters, -1, &c->dt, &c->it, &c->t), "adv_run_until");
#endif
    die_on(adv_sync(s), "adv_sync");
    pthread_barrier_wait(c->bar);
    c->compute = now_s() - t0;

    t0 = now_s();
    die_on(adv_download(s, h_s_u), "adv_download");
    c->DtH = now_s() - t0;
    pthread_barrier_wait(c->bar);
    /* Merge_domains (Tools.c:205-224): owned planes only */
    memcpy(c->h_u + xy * ((size_t)c->d.rank * c->_Nz + RADIUS), h_s_u + xy * RADIUS, sizeof(double) * xy * c->_Nz);
    adv_destroy(s);
    free(h_s_u);
    return NULL;
}

int main(int argc, char** argv)
{
    double a1, a2 = 0, L, W, H;
    unsigned Nx, Ny, Nz, iters = 0;
    (void)iters;
    if (argc == 12) {
#ifdef BURGERS
        a1 = atof(argv[1]); a2 = atof(argv[2]); L = atof(argv[3]); W = atof(argv[4]); H = atof(argv[5]);
        Nx = atoi(argv[6]); Ny = atoi(argv[7]); Nz = atoi(argv[8]);
#else
        a1 = atof(argv[1]); L = atof(argv[2]); W = atof(argv[3]); H = atof(argv[4]);
        Nx = atoi(argv[5]); Ny = atoi(argv[6]); Nz = atoi(argv[7]); iters = atoi(argv[8]);
#endif
    } else {
#ifdef BURGERS
        printf("Usage: %s tEnd CFL L W H Nx Ny NZ block_x block_y block_z\n", argv[0]);
#else
        printf("Usage: %s K L W H Nx Ny NZ max_iters block_x block_y block_z\n", argv[0]);
#endif
        exit(1);
    }
    int P = getenv("ADV_NGPUS") ? atoi(getenv("ADV_NGPUS")) : 1, ndev = 0;
    die_on(adv_device_count(&ndev), "adv_device_count");
    if (P < 1 || P > ndev) { printf("ADV_NGPUS=%d but %d B200 device(s) visible\n", P, ndev); exit(-1); }

    const double dx = L / (Nx - 1), dy = W / (Ny - 1), dz = H / (Nz - 1);
    const unsigned NZ = Nz + 2 * RADIUS;
    unsigned char id[128];
    memset(id, 0, sizeof(id));
    /* no loadable NCCL on this host: the ranks still exchange halos through peer stores (the default below) */
    if (P > 1 && adv_comm_unique_id(id)) printf("NCCL unavailable (%s): peer-store halo exchange only\n", adv_last_error());
    double* h_u = (double*)malloc(sizeof(double) * Nx * Ny * NZ);
    pthread_barrier_t bar;
    pthread_barrier_init(&bar, NULL, P);
    rank_ctx* ctx = (rank_ctx*)calloc(P, sizeof(rank_ctx));
    pthread_t* th = (pthread_t*)calloc(P, sizeof(pthread_t));
    for (int r = 0; r < P; r++) {
        adv_desc* d = &ctx[r].d;
        memset(d, 0, sizeof(*d));
#ifdef BURGERS
        d->target = ADV_TARGET_BURGERS3D_MGPU; d->coef = a2; ctx[r].tEnd = a1;
#else
        d->target = ADV_TARGET_HEAT3D_MGPU; d->coef = a1; ctx[r].iters = iters;
#endif
        d->precision = ADV_F64; d->nx = Nx; d->ny = Ny; d->nz = Nz; d->lx = L; d->ly = W; d->lz = H;
        d->device = r; d->nranks = P; d->rank = r;
        memcpy(d->comm_id, id, 128);
        ctx[r].h_u = h_u; ctx[r]._Nz = Nz / P; ctx[r].bar = &bar; ctx[r].all = ctx;
    }
#ifdef BURGERS
    printf("dx: %g, dy: %g, dz: %g, final time: %g\n\n", dx, dy, dz, a1);
#else
    {
        const double dt = 1 / (2 * a1 * (1 / dx / dx + 1 / dy / dy + 1 / dz / dz)) * 0.8;
        printf("dx: %g, dy: %g, dz: %g, final time: %g\n\n", dx, dy, dz, dt * iters);
    }
#endif
    for (int r = 0; r < P; r++) pthread_create(&th[r], NULL, rank_main, &ctx[r]);
    for (int r = 0; r < P; r++) pthread_join(th[r], NULL);

#ifdef BURGERS
    printf("dt: %g, iterations: %g, final time: %g\n\n", ctx[0].dt, ctx[0].it, ctx[0].t);
#else
    printf("dt: %g, iterations: %d, final time: %g\n\n", ctx[0].dt, (int)ctx[0].it, ctx[0].t);
#endif
    save_binary_f32(h_u, (size_t)Nx * Ny * NZ, "result.bin");
    float gflops = calc_gflops_stages((float)ctx[0].compute, (unsigned)ctx[0].it, Nx, Ny, NZ);
#ifdef BURGERS
    print_summary_mgpu("=======================", "Burgers-3D MPI-GPU-WENO5", "=====================", "Pitched Memory", ctx[0].compute, ctx[0].HtD, ctx[0].DtH, gflops, (int)ctx[0].it, Nx, Ny, NZ);
#else
    print_summary_mgpu("=======================", "Diffusion-3D MPI-GPU-FD4", "=====================", "Pitched Memory", ctx[0].compute, ctx[0].HtD, ctx[0].DtH, gflops, (int)ctx[0].it, Nx, Ny, NZ);
#endif
    print_b200_lines(ctx[0].compute, ctx[0].it, (double)Nx * Ny * Nz, 8);
    free(h_u); free(ctx); free(th);
    return 0;
}
