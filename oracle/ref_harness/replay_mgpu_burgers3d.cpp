/* TEST INFRASTRUCTURE ONLY (part of oracle/_ref).
 * Replays MultiGPU/Burgers3d_Baseline/main.c:60-349 on the CPU for P emulated
 * MPI ranks inside one process, calling the reference's own kernel bodies
 * (piped in above by oracle/Makefile) and its own Tools.c (Init_domain,
 * Init_subdomain, Merge_domains, SaveBinary3D).  MPI messages become memcpy
 * between per-rank buffers; every rank's enqueue order is kept (right slab,
 * left slab, interior, unpack, RK), ranks are visited 0..P-1 in two phases
 * (compute+pack for all, then unpack+RK for all), i.e. the order the author's
 * stream/MPI synchronisation intends.  pitch = Nx, gc_pitch = Nx.
 */
static int ref_getBlock(int n, int block)   /* MultiGPU/Burgers3d_Baseline/Util.cu:98-101 */
{
    return (n+2)/block + ((n+2)%block == 0 ? 0 : 1);
}

extern "C" int ref_init(int IC, unsigned Nx, unsigned Ny, unsigned Nz, double L, double W, double H, double* h_u)
{
    const REAL dx = L/(Nx-1), dy = W/(Ny-1), dz = H/(Nz-1);
    Init_domain(IC, h_u, dx, dy, dz, Nx, Ny, Nz+2*RADIUS);
    return 0;
}

extern "C" int ref_save(double* u, unsigned Nx, unsigned Ny, unsigned NZ, const char* name)
{
    SaveBinary3D(u, Nx, Ny, NZ, name);
    return 0;
}

extern "C" int ref_run(int numberOfProcesses, double tEnd, double CFL, double L, double W, double H,
                       unsigned Nx, unsigned Ny, unsigned Nz,
                       unsigned blockX, unsigned blockY, unsigned blockZ,
                       const double* u_in, double* u_out,
                       double* dt_out, double* it_out, double* t_out)
{
    const int P = numberOfProcesses;
    const REAL dx = L/(Nx-1);
    const REAL dy = W/(Ny-1);
    const REAL dz = H/(Nz-1);
    const unsigned _Nz = Nz/P;
    const unsigned  NZ = Nz+2*RADIUS;
    const unsigned _NZ = _Nz+2*RADIUS;
    const unsigned pitch = Nx, gc_pitch = Nx;
    const size_t ns = (size_t)Nx*Ny*_NZ, nh = (size_t)Nx*Ny*RADIUS;

    REAL* h_u = (REAL*)malloc(sizeof(REAL)*Nx*Ny*NZ);
    memcpy(h_u, u_in, sizeof(REAL)*Nx*Ny*NZ);

    REAL **d_s_u = new REAL*[P], **d_s_uo = new REAL*[P], **d_s_Lu = new REAL*[P];
    REAL **l_send = new REAL*[P], **r_send = new REAL*[P], **l_recv = new REAL*[P], **r_recv = new REAL*[P];
    for (int r = 0; r < P; r++) {
        d_s_u[r]  = (REAL*)malloc(sizeof(REAL)*ns);
        d_s_uo[r] = (REAL*)malloc(sizeof(REAL)*ns);
        d_s_Lu[r] = (REAL*)malloc(sizeof(REAL)*ns);
        l_send[r] = (REAL*)malloc(sizeof(REAL)*nh); r_send[r] = (REAL*)malloc(sizeof(REAL)*nh);
        l_recv[r] = (REAL*)malloc(sizeof(REAL)*nh); r_recv[r] = (REAL*)malloc(sizeof(REAL)*nh);
        Init_subdomain(h_u, d_s_u[r], r, Nx, Ny, _Nz);
    }

    dim3 threadsPerBlock3D(blockX, blockY, blockZ);
    unsigned blocksInX = ref_getBlock( Nx, blockX);
    unsigned blocksInY = ref_getBlock( Ny, blockY);
    unsigned blocksInZ = ref_getBlock(_NZ, blockZ);
    dim3 numBlocks3D(blocksInX, blocksInY, blocksInZ);
    dim3 threadsPerHalo2D_XY(blockX,blockY,1);
    dim3 numBlocksHalo2D_XY(blocksInX,blocksInY,1);
    dim3 threadsPerBlock2D_XY(blockX,blockY,1); dim3 numBlocks2D_XY(blocksInX,blocksInY,1);
    dim3 threadsPerBlock2D_XZ(blockX,1,blockZ); dim3 numBlocks2D_XZ(blocksInX,1,blocksInZ);
    dim3 threadsPerBlock2D_YZ(1,blockY,blockZ); dim3 numBlocks2D_YZ(1,blocksInY,blocksInZ);

#define ADV(r, kstart, kstop) do { \
    LAUNCH(numBlocks2D_YZ, threadsPerBlock2D_YZ, Compute_dF(d_s_u[r], d_s_Lu[r], pitch, Nx, Ny, _NZ, kstart, kstop, LOOP, dx)); \
    LAUNCH(numBlocks2D_XZ, threadsPerBlock2D_XZ, Compute_dG(d_s_u[r], d_s_Lu[r], pitch, Nx, Ny, _NZ, kstart, kstop, LOOP, dy)); \
    LAUNCH(numBlocks2D_XY, threadsPerBlock2D_XY, Compute_dH(d_s_u[r], d_s_Lu[r], pitch, Nx, Ny, _NZ, kstart, kstop, LOOP, dz)); \
  } while (0)

    REAL dt = 0, it = 0, t = 0;
    while (t < tEnd) {
        dt = CFL*dx/1.0; if ((t+dt) > tEnd) { dt = tEnd-t; }
        t += dt; it += 1;
        for (int r = 0; r < P; r++) {
            memcpy(d_s_uo[r], d_s_u[r], sizeof(REAL)*ns);
            memset(d_s_Lu[r], 0, sizeof(REAL)*ns);
        }
        for (unsigned step = 1; step <= 3; step++) {
            for (int rank = 0; rank < P; rank++) {
                if (rank < P-1) {
                    unsigned kstart = _Nz, kstop = _Nz+RADIUS;
                    ADV(rank, kstart, kstop);
                    LAUNCH(numBlocksHalo2D_XY, threadsPerHalo2D_XY,
                           copy_br_to_gc(d_s_Lu[rank], r_send[rank], Nx, Ny, _NZ, pitch, gc_pitch, 0));
                }
                if (rank > 0) {
                    unsigned kstart = RADIUS, kstop = 2*RADIUS;
                    ADV(rank, kstart, kstop);
                    LAUNCH(numBlocksHalo2D_XY, threadsPerHalo2D_XY,
                           copy_br_to_gc(d_s_Lu[rank], l_send[rank], Nx, Ny, _NZ, pitch, gc_pitch, 1));
                }
                if (rank == 0) { unsigned kstart = RADIUS, kstop = _Nz; ADV(rank, kstart, kstop); }
                if (rank > 0 && rank < P-1) { unsigned kstart = 2*RADIUS, kstop = _Nz; ADV(rank, kstart, kstop); }
                if (rank == P-1) { unsigned kstart = 2*RADIUS, kstop = _Nz+RADIUS; ADV(rank, kstart, kstop); }
            }
            for (int rank = 0; rank < P; rank++) {
                if (rank < P-1) {   /* tag 5: what rank+1 sent left */
                    memcpy(r_recv[rank], l_send[rank+1], sizeof(REAL)*nh);
                    LAUNCH(numBlocksHalo2D_XY, threadsPerHalo2D_XY,
                           copy_gc_to_br(d_s_Lu[rank], r_recv[rank], Nx, Ny, _NZ, pitch, gc_pitch, 0));
                }
                if (rank > 0) {     /* tag 1: what rank-1 sent right */
                    memcpy(l_recv[rank], r_send[rank-1], sizeof(REAL)*nh);
                    LAUNCH(numBlocksHalo2D_XY, threadsPerHalo2D_XY,
                           copy_gc_to_br(d_s_Lu[rank], l_recv[rank], Nx, Ny, _NZ, pitch, gc_pitch, 1));
                }
                LAUNCH(numBlocks3D, threadsPerBlock3D,
                       Compute_RK(d_s_u[rank], d_s_uo[rank], d_s_Lu[rank], step, pitch, Nx, Ny, _NZ, dt));
            }
        }
    }
    for (int r = 0; r < P; r++) Merge_domains(d_s_u[r], h_u, r, Nx, Ny, _Nz);
    memcpy(u_out, h_u, sizeof(REAL)*Nx*Ny*NZ);
    if (dt_out) *dt_out = dt;
    if (it_out) *it_out = it;
    if (t_out)  *t_out  = t;
    for (int r = 0; r < P; r++) {
        free(d_s_u[r]); free(d_s_uo[r]); free(d_s_Lu[r]);
        free(l_send[r]); free(r_send[r]); free(l_recv[r]); free(r_recv[r]);
    }
    delete[] d_s_u; delete[] d_s_uo; delete[] d_s_Lu;
    delete[] l_send; delete[] r_send; delete[] l_recv; delete[] r_recv;
    free(h_u);
    return 0;
}
