/* TEST INFRASTRUCTURE ONLY (part of oracle/_ref).
 * Replays SingleGPU/Diffusion2d/main.cpp:53-125 on the CPU with the
 * reference's own kernel bodies (piped in above by oracle/Makefile).
 * The reference's Compute_Laplace2d writes Lu=0 for threads outside the grid
 * without a bounds guard (kernels.cu:86-87); the harness therefore only
 * accepts Nx%blockX==0 and Ny%blockY==0, where no such thread exists.
 */
extern "C" int ref_init(unsigned Nx, unsigned Ny, unsigned L, unsigned W, double* u)
{
    REAL dx = (REAL)L/(Nx-1), dy = (REAL)W/(Ny-1);
    Call_Init2d(3, u, dx, dy, Nx, Ny);
    return 0;
}

extern "C" int ref_run(double C, unsigned L, unsigned W, unsigned Nx, unsigned Ny,
                       unsigned max_iters, unsigned blockX, unsigned blockY,
                       const double* u_in, double* u_out, double* dt_out)
{
    if (Nx % blockX || Ny % blockY) return 1;
    REAL dx, dy, dt, kx, ky;
    dx = (REAL)L/(Nx-1);
    dy = (REAL)W/(Ny-1);
    dt = 1/(2*C*(1/dx/dx+1/dy/dy))*0.9;
    kx = C/(12*dx*dx);
    ky = C/(12*dy*dy);
    d_kx = kx; d_ky = ky;
    if (dt_out) *dt_out = dt;

    size_t n = (size_t)Nx*Ny;
    REAL* d_u  = (REAL*)malloc(sizeof(REAL)*n);
    REAL* d_uo = (REAL*)malloc(sizeof(REAL)*n);
    REAL* d_Lu = (REAL*)malloc(sizeof(REAL)*n);
    memcpy(d_u, u_in, sizeof(REAL)*n);

    dim3 threadsPerBlock2D(blockX,blockY,1);
    dim3 numBlocks2D(DIVIDE_INTO(Nx,blockX), DIVIDE_INTO(Ny,blockY), 1);

    for (unsigned it = 0; it < max_iters; it++) {
        memcpy(d_uo, d_u, sizeof(REAL)*n);
        memset(d_Lu, 0, sizeof(REAL)*n);
        for (unsigned step = 1; step <= 3; step++) {
            LAUNCH(numBlocks2D, threadsPerBlock2D, Compute_Laplace2d(d_u, d_Lu, Nx, Ny));
            LAUNCH(numBlocks2D, threadsPerBlock2D, Compute_RK(d_u, d_uo, d_Lu, step, Nx, Ny, dt));
        }
    }
    memcpy(u_out, d_u, sizeof(REAL)*n);
    free(d_u); free(d_uo); free(d_Lu);
    return 0;
}
