/* TEST INFRASTRUCTURE ONLY (part of oracle/_ref).
 * Replays MultiGPU/Diffusion2d_Baseline/main.c:57-320 on the CPU for P emulated MPI ranks, calling the
 * reference's own kernel bodies (piped in above by oracle/Makefile) and its own Tools.c (built with
 * -DHALO=6, the macro Init_subdomain uses but the header never defines).  Two-phase rank order as in the
 * 3-D replays (compute + pack for all ranks, then unpack + RK for all).  pitch = Nx, gc_pitch = Nx.
 * Launch geometry as in main.c:147-156: blocks DIVIDE_INTO(Nx,bx) x DIVIDE_INTO(_NY,by); the boundary
 * launches use ONE block row, so they cover rows [jstart-1, jstart-1+by) -- with the shipped 32x32 block
 * (run.sh) that is the whole [jstart-1, jstop) window.
 */
extern "C" int ref_init(int IC, unsigned Nx, unsigned Ny, double L, double W, double* h_u)
{
    const REAL dx = L/(Nx-1), dy = W/(Ny-1);
    Init_domain(IC, h_u, dx, dy, Nx, Ny+2*RADIUS);
    return 0;
}

extern "C" int ref_save(double* u, unsigned Nx, unsigned NY, const char* name)
{
    SaveBinary2D(u, Nx, NY, name);
    return 0;
}

extern "C" int ref_run(int numberOfProcesses, double K, double L, double W,
                       unsigned Nx, unsigned Ny, unsigned max_iters, unsigned blockX, unsigned blockY,
                       const double* u_in, double* u_out, double* dt_out, int* it_out, double* t_out)
{
    const int P = numberOfProcesses;
    const REAL dx = L/(Nx-1);
    const REAL dy = W/(Ny-1);
    const REAL dt = 1/(2*K*(1/dx/dx+1/dy/dy))*0.8;
    const REAL kx = K/(12*dx*dx);
    const REAL ky = K/(12*dy*dy);
    const REAL tEnd = dt*max_iters;
    const unsigned _Ny = Ny/P;
    const unsigned  NY = Ny+2*RADIUS;
    const unsigned _NY = _Ny+2*RADIUS;
    const unsigned pitch = Nx, gc_pitch = Nx;
    const size_t ns = (size_t)Nx*_NY, nh = (size_t)Nx*RADIUS;

    REAL* h_u = (REAL*)malloc(sizeof(REAL)*Nx*NY);
    memcpy(h_u, u_in, sizeof(REAL)*Nx*NY);

    REAL **d_s_u = new REAL*[P], **d_s_uo = new REAL*[P], **d_s_Lu = new REAL*[P];
    REAL **l_send = new REAL*[P], **r_send = new REAL*[P], **l_recv = new REAL*[P], **r_recv = new REAL*[P];
    for (int r = 0; r < P; r++) {
        d_s_u[r]  = (REAL*)malloc(sizeof(REAL)*ns);
        d_s_uo[r] = (REAL*)malloc(sizeof(REAL)*ns);
        d_s_Lu[r] = (REAL*)calloc(ns, sizeof(REAL));    /* memset once, main.c:175 */
        l_send[r] = (REAL*)malloc(sizeof(REAL)*nh); r_send[r] = (REAL*)malloc(sizeof(REAL)*nh);
        l_recv[r] = (REAL*)malloc(sizeof(REAL)*nh); r_recv[r] = (REAL*)malloc(sizeof(REAL)*nh);
        Init_subdomain(h_u, d_s_u[r], r, Nx, _Ny);
    }

    dim3 threadsPerBlock2D(blockX, blockY);
    unsigned blocksInX = DIVIDE_INTO( Nx,blockX);
    unsigned blocksInY = DIVIDE_INTO(_NY,blockY);
    dim3 numBlocks2D(blocksInX, blocksInY);
    dim3 threadsPerHalo1D(blockX);
    dim3 numBlocksHalo2D(blocksInX,1);
    dim3 numBlocksHalo1D(blocksInX);

#define DIFF(grid, r, jstart, jstop) \
    LAUNCH(grid, threadsPerBlock2D, Compute_Laplace(d_s_u[r], d_s_Lu[r], kx, ky, pitch, Nx, _NY, jstart, jstop))

    int it = 0; REAL t = 0;
    while (t < tEnd) {
        t += dt; it += 1;
        for (int r = 0; r < P; r++) memcpy(d_s_uo[r], d_s_u[r], sizeof(REAL)*ns);
        for (unsigned step = 1; step <= 3; step++) {
            for (int rank = 0; rank < P; rank++) {
                if (rank < P-1) {
                    unsigned jstart = _Ny, jstop = _Ny+RADIUS;
                    DIFF(numBlocksHalo2D, rank, jstart, jstop);
                    LAUNCH(numBlocksHalo1D, threadsPerHalo1D,
                           copy_br_to_gc(d_s_Lu[rank], r_send[rank], Nx, _NY, pitch, gc_pitch, 0));
                }
                if (rank > 0) {
                    unsigned jstart = RADIUS, jstop = 2*RADIUS;
                    DIFF(numBlocksHalo2D, rank, jstart, jstop);
                    LAUNCH(numBlocksHalo1D, threadsPerHalo1D,
                           copy_br_to_gc(d_s_Lu[rank], l_send[rank], Nx, _NY, pitch, gc_pitch, 1));
                }
                if (rank == 0) { unsigned jstart = RADIUS, jstop = _Ny; DIFF(numBlocks2D, rank, jstart, jstop); }
                if (rank > 0 && rank < P-1) { unsigned jstart = 2*RADIUS, jstop = _Ny; DIFF(numBlocks2D, rank, jstart, jstop); }
                if (rank == P-1) { unsigned jstart = 2*RADIUS, jstop = _Ny+RADIUS; DIFF(numBlocks2D, rank, jstart, jstop); }
            }
            for (int rank = 0; rank < P; rank++) {
                if (rank < P-1) {
                    memcpy(r_recv[rank], l_send[rank+1], sizeof(REAL)*nh);
                    LAUNCH(numBlocksHalo1D, threadsPerHalo1D,
                           copy_gc_to_br(d_s_Lu[rank], r_recv[rank], Nx, _NY, pitch, gc_pitch, 0));
                }
                if (rank > 0) {
                    memcpy(l_recv[rank], r_send[rank-1], sizeof(REAL)*nh);
                    LAUNCH(numBlocksHalo1D, threadsPerHalo1D,
                           copy_gc_to_br(d_s_Lu[rank], l_recv[rank], Nx, _NY, pitch, gc_pitch, 1));
                }
                LAUNCH(numBlocks2D, threadsPerBlock2D,
                       Compute_RK(d_s_u[rank], d_s_uo[rank], d_s_Lu[rank], step, pitch, Nx, _NY, dt));
            }
        }
    }
    for (int r = 0; r < P; r++) Merge_domains(d_s_u[r], h_u, r, Nx, _Ny);
    memcpy(u_out, h_u, sizeof(REAL)*Nx*NY);
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
