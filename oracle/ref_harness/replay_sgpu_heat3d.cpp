/* TEST INFRASTRUCTURE ONLY (part of oracle/_ref).
 * Replays the host loop of the reference's single-GPU 3-D heat drivers on the
 * CPU, calling the reference's own kernel bodies (piped in above this text by
 * oracle/Makefile).  -DREF_BLOCKING=0: SingleGPU/Diffusion3d_baselineCode
 * (main.cpp:55-121); -DREF_BLOCKING=1: SingleGPU/Diffusion3d_Blocking
 * (main.cpp:50-121, launch shapes :87-98, pitch taken as Nx).
 */
extern "C" int ref_init(unsigned Nx, unsigned Ny, unsigned Nz, unsigned L, unsigned W, unsigned H, double* u)
{
    REAL dx = (REAL)L/(Nx-1), dy = (REAL)W/(Ny-1), dz = (REAL)H/(Nz-1);
    Call_Init3D(3, u, dx, dy, dz, Nx, Ny, Nz);
    return 0;
}

extern "C" int ref_run(double C, unsigned L, unsigned W, unsigned H,
                       unsigned Nx, unsigned Ny, unsigned Nz, unsigned max_iters,
                       unsigned blockX, unsigned blockY, unsigned blockZ,
                       const double* u_in, double* u_out, double* dt_out)
{
    REAL dx, dy, dz, dt, kx, ky, kz;
    dx = (REAL)L/(Nx-1);
    dy = (REAL)W/(Ny-1);
    dz = (REAL)H/(Nz-1);
    dt = 1/(2*C*(1/dx/dx+1/dy/dy+1/dz/dz))*0.9;
    kx = C/(12*dx*dx);
    ky = C/(12*dy*dy);
    kz = C/(12*dz*dz);
    d_kx = kx; d_ky = ky; d_kz = kz;       /* stands in for CopyToConstantMemory */
    if (dt_out) *dt_out = dt;

    size_t n = (size_t)Nx*Ny*Nz;
    REAL* d_u  = (REAL*)malloc(sizeof(REAL)*n);
    REAL* d_uo = (REAL*)malloc(sizeof(REAL)*n);
    REAL* d_Lu = (REAL*)calloc(n, sizeof(REAL));   /* cudaMemset once */
    memcpy(d_u, u_in, sizeof(REAL)*n);

#if REF_BLOCKING
    (void)blockX; (void)blockY; (void)blockZ;
    unsigned pitch = Nx;
    dim3 threadsPerBlock3D(32,4,4);
    dim3 numBlocks3D(DIVIDE_INTO(Nx,32), DIVIDE_INTO(Ny,4), DIVIDE_INTO(Nz,4));
    dim3 threadsPerBlock2D(32,4,1);
    dim3 numBlocks2D(DIVIDE_INTO(Nx,32), DIVIDE_INTO(Ny,4), DIVIDE_INTO(Nz,k_loop));
#else
    dim3 threadsPerBlock3D(blockX,blockY,blockZ);
    dim3 numBlocks3D(DIVIDE_INTO(Nx,blockX), DIVIDE_INTO(Ny,blockY), DIVIDE_INTO(Nz,blockZ));
#endif

    for (unsigned it = 0; it < max_iters; it++) {
        memcpy(d_uo, d_u, sizeof(REAL)*n);
        for (unsigned step = 1; step <= 3; step++) {
#if REF_BLOCKING
            LAUNCH(numBlocks2D, threadsPerBlock2D,
                   Compute_Laplace3d_Async(d_u, d_Lu, pitch, Nx, Ny, Nz, 3, Nz-2, k_loop));
            LAUNCH(numBlocks3D, threadsPerBlock3D,
                   Compute_RK(d_u, d_uo, d_Lu, step, pitch, Nx, Ny, Nz, dt));
#else
            LAUNCH(numBlocks3D, threadsPerBlock3D, Compute_Laplace3d(d_u, d_Lu, Nx, Ny, Nz));
            LAUNCH(numBlocks3D, threadsPerBlock3D, Compute_RK(d_u, d_uo, d_Lu, step, Nx, Ny, Nz, dt));
#endif
        }
    }
    memcpy(u_out, d_u, sizeof(REAL)*n);
    free(d_u); free(d_uo); free(d_Lu);
    return 0;
}
