/* TEST INFRASTRUCTURE ONLY (part of oracle/_ref).
 * Replays MultiGPU/Diffusion3d_Baseline/main.c:60-336 on the CPU for P
 * emulated MPI ranks, calling the reference's own kernel bodies (piped in
 * above by oracle/Makefile) and its own Tools.c (built with -DHALO=6, the
 * macro that file uses but never defines).  Same two-phase rank order as the
 * Burgers replay.  pitch = Nx, gc_pitch = Nx.
 * The reference's LaplaceO4_async writes the first plane of every z-block
 * without a bounds check (Kernels.cu:232-243); callers must pick sizes where
 * that plane lies inside the slab (see tests), as the reference's own runs do.
 */
static int ref_getBlock(int n, int block)   /* MultiGPU/Diffusion3d_Baseline/Util.cu:98-101 */
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

extern "C" int ref_run(int numberOfProcesses, double K, double L, double W, double H,
                       unsigned Nx, unsigned Ny, unsigned Nz, unsigned max_iters,
                       unsigned blockX, unsigned blockY, unsigned blockZ,
                       const double* u_in, double* u_out,
                       double* dt_out, int* it_out, double* t_out)
{
    const int P = numberOfProcesses;
    const REAL dx = L/(Nx-1);
    const REAL dy = W/(Ny-1);
    const REAL dz = H/(Nz-1);
    const REAL dt = 1/(2*K*(1/dx/dx+1/dy/dy+1/dz/dz))*0.8;
    const REAL kx = K/(12*dx*dx);
    const REAL ky = K/(12*dy*dy);
    const REAL kz = K/(12*dz*dz);
    const REAL tEnd = dt*max_iters;
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
        d_s_Lu[r] = (REAL*)calloc(ns, sizeof(REAL));    /* memset once, main.c:182 */
        l_send[r] = (REAL*)malloc(sizeof(REAL)*nh); r_send[r] = (REAL*)malloc(sizeof(REAL)*nh);
        l_recv[r] = (REAL*)malloc(sizeof(REAL)*nh); r_recv[r] = (REAL*)malloc(sizeof(REAL)*nh);
        Init_subdomain(h_u, d_s_u[r], r, Nx, Ny, _Nz);
    }

    dim3 threadsPerBlock3D(blockX, blockY, blockZ);
    unsigned blocksInX = ref_getBlock( Nx, blockX);
    unsigned blocksInY = ref_getBlock( Ny, blockY);
    unsigned blocksInZ = ref_getBlock(_NZ,  LOOP );
    dim3 numBlocks3D(blocksInX, blocksInY, blocksInZ);
    dim3 numBlocksHalo2D_XY(blocksInX,blocksInY,1);
    dim3 numBlocksHalo2D_XY_RADIUS(blocksInX,blocksInY,RADIUS);
    dim3 threadsPerBlock3D_RK(8,8,8);
    dim3 numBlocks3D_RK(ref_getBlock(Nx,8), ref_getBlock(Ny,8), ref_getBlock(_NZ,8));

#define DIFF(grid, r, kstart, kstop) \
    LAUNCH(grid, threadsPerBlock3D, LaplaceO4_async(d_s_u[r], d_s_Lu[r], kx, ky, kz, pitch, Nx, Ny, _NZ, kstart, kstop, LOOP))

    int it = 0; REAL t = 0;
    while (t < tEnd) {
        t += dt; it += 1;
        for (int r = 0; r < P; r++) memcpy(d_s_uo[r], d_s_u[r], sizeof(REAL)*ns);
        for (unsigned step = 1; step <= 3; step++) {
            for (int rank = 0; rank < P; rank++) {
                if (rank < P-1) {
                    unsigned kstart = _Nz, kstop = _Nz+RADIUS;
                    DIFF(numBlocksHalo2D_XY_RADIUS, rank, kstart, kstop);
                    LAUNCH(numBlocksHalo2D_XY, threadsPerBlock3D,
                           copy_br_to_gc(d_s_Lu[rank], r_send[rank], Nx, Ny, _NZ, pitch, gc_pitch, 0));
                }
                if (rank > 0) {
                    unsigned kstart = RADIUS, kstop = 2*RADIUS;
                    DIFF(numBlocksHalo2D_XY_RADIUS, rank, kstart, kstop);
                    LAUNCH(numBlocksHalo2D_XY, threadsPerBlock3D,
                           copy_br_to_gc(d_s_Lu[rank], l_send[rank], Nx, Ny, _NZ, pitch, gc_pitch, 1));
                }
                if (rank == 0) { unsigned kstart = RADIUS, kstop = _Nz; DIFF(numBlocks3D, rank, kstart, kstop); }
                if (rank > 0 && rank < P-1) { unsigned kstart = 2*RADIUS, kstop = _Nz; DIFF(numBlocks3D, rank, kstart, kstop); }
                if (rank == P-1) { unsigned kstart = 2*RADIUS, kstop = _Nz+RADIUS; DIFF(numBlocks3D, rank, kstart, kstop); }
            }
            for (int rank = 0; rank < P; rank++) {
                if (rank < P-1) {
                    memcpy(r_recv[rank], l_send[rank+1], sizeof(REAL)*nh);
                    LAUNCH(numBlocksHalo2D_XY, threadsPerBlock3D,
                           copy_gc_to_br(d_s_Lu[rank], r_recv[rank], Nx, Ny, _NZ, pitch, gc_pitch, 0));
                }
                if (rank > 0) {
                    memcpy(l_recv[rank], r_send[rank-1], sizeof(REAL)*nh);
                    LAUNCH(numBlocksHalo2D_XY, threadsPerBlock3D,
                           copy_gc_to_br(d_s_Lu[rank], l_recv[rank], Nx, Ny, _NZ, pitch, gc_pitch, 1));
                }
                LAUNCH(numBlocks3D_RK, threadsPerBlock3D_RK,
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
