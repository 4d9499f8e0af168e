/* TEST INFRASTRUCTURE ONLY (part of oracle/_ref).
 * Replays MultiGPU/Burgers2d_Baseline/main.c:57-330 on the CPU for P emulated MPI ranks, calling the
 * reference's own kernel bodies (piped in above by oracle/Makefile) and its own Tools.c.  Two-phase rank
 * order as in the 3-D replays.  pitch = Nx, gc_pitch = Nx.  Launch geometry as in main.c:141-161: the
 * x sweep runs one thread per row (grid (1, DIVIDE_INTO(_NY,by)), block (1,by)), the y sweep one thread
 * per column (grid (DIVIDE_INTO(Nx,bx),1), block (bx,1)).
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

extern "C" int ref_run(int numberOfProcesses, double tEnd_, double CFL_, double L, double W,
                       unsigned Nx, unsigned Ny, unsigned blockX, unsigned blockY,
                       const double* u_in, double* u_out, double* dt_out, double* it_out, double* t_out)
{
    const int P = numberOfProcesses;
    const REAL tEnd = tEnd_, CFL = CFL_;
    REAL dx = L/(Nx-1);
    REAL dy = W/(Ny-1);
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
        d_s_Lu[r] = (REAL*)calloc(ns, sizeof(REAL));    /* memset once, main.c:172 */
        l_send[r] = (REAL*)malloc(sizeof(REAL)*nh); r_send[r] = (REAL*)malloc(sizeof(REAL)*nh);
        l_recv[r] = (REAL*)malloc(sizeof(REAL)*nh); r_recv[r] = (REAL*)malloc(sizeof(REAL)*nh);
        Init_subdomain(h_u, d_s_u[r], r, Nx, _Ny);
    }

    unsigned blocksInX = DIVIDE_INTO( Nx,blockX);
    unsigned blocksInY = DIVIDE_INTO(_NY,blockY);
    dim3 threadsPerBlock2D(blockX, blockY);
    dim3 numBlocks2D(blocksInX, blocksInY);
    dim3 threadsPerHalo1D(blockX);
    dim3 numBlocksHalo1D(blocksInX);
    dim3 threadsPerBlock_Y(1,blockY);
    dim3 numBlocks_Y(1,blocksInY);
    dim3 threadsPerBlock_X(blockX,1);
    dim3 numBlocks_X(blocksInX,1);

#define ADV(r, jstart, jstop) do { \
    LAUNCH(numBlocks_Y, threadsPerBlock_Y, Compute_dF(d_s_u[r], d_s_Lu[r], pitch, Nx, _NY, jstart, jstop, dx)); \
    LAUNCH(numBlocks_X, threadsPerBlock_X, Compute_dG(d_s_u[r], d_s_Lu[r], pitch, Nx, _NY, jstart, jstop, dy)); \
} while (0)

    REAL dt = 0, it = 0, t = 0;
    while (t < tEnd) {
        dt = CFL*dx/1.0; if ((t+dt) > tEnd) { dt = tEnd-t; }
        t += dt; it += 1;
        for (int r = 0; r < P; r++) memcpy(d_s_uo[r], d_s_u[r], sizeof(REAL)*ns);
        for (unsigned step = 1; step <= 3; step++) {
            for (int rank = 0; rank < P; rank++) {
                if (rank < P-1) {
                    unsigned jstart = _Ny, jstop = _Ny+RADIUS;
                    ADV(rank, jstart, jstop);
                    LAUNCH(numBlocksHalo1D, threadsPerHalo1D,
                           copy_br_to_gc(d_s_Lu[rank], r_send[rank], Nx, _NY, pitch, gc_pitch, 0));
                }
                if (rank > 0) {
                    unsigned jstart = RADIUS, jstop = 2*RADIUS;
                    ADV(rank, jstart, jstop);
                    LAUNCH(numBlocksHalo1D, threadsPerHalo1D,
                           copy_br_to_gc(d_s_Lu[rank], l_send[rank], Nx, _NY, pitch, gc_pitch, 1));
                }
                if (rank == 0) { unsigned jstart = RADIUS, jstop = _Ny; ADV(rank, jstart, jstop); }
                if (rank > 0 && rank < P-1) { unsigned jstart = 2*RADIUS, jstop = _Ny; ADV(rank, jstart, jstop); }
                if (rank == P-1) { unsigned jstart = 2*RADIUS, jstop = _Ny+RADIUS; ADV(rank, jstart, jstop); }
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
