/* TEST INFRASTRUCTURE ONLY (part of oracle/_ref).
 * Replays SingleGPU/Burgers3d_WENO5/main.cpp:52-149 on the CPU with the
 * reference's own kernel bodies (piped in above by oracle/Makefile):
 * WENO5-Z + Lax-Friedrichs split, C=1e-5 viscous 13-point term, SSP-RK3.
 * laplace_mode 0: Compute_Laplace_Async exactly as shipped (kernels.cu:564;
 *   consecutive z-blocks overlap, so planes 16..18 receive the viscous term
 *   twice when executed in program order);
 * laplace_mode 1: the reference's non-blocked Compute_Laplace (kernels.cu:399-
 *   439, commented out at :563) -- each plane exactly once.
 */
extern "C" int ref_init(unsigned Nx, unsigned Ny, unsigned Nz, double L, double W, double H, double* u)
{
    REAL dx = (REAL)L/(Nx-1), dy = (REAL)W/(Ny-1), dz = (REAL)H/(Nz-1);
    Call_Init3d(3, u, dx, dy, dz, Nx, Ny, Nz);
    return 0;
}

extern "C" int ref_run(double tEnd, double CFL, double L, double W, double H,
                       unsigned Nx, unsigned Ny, unsigned Nz,
                       unsigned blockX, unsigned blockY, unsigned blockZ, int laplace_mode,
                       const double* u_in, double* u_out,
                       double* dt_out, double* it_out, double* t_out)
{
    REAL dx = (REAL)L/(Nx-1);
    REAL dy = (REAL)W/(Ny-1);
    REAL dz = (REAL)H/(Nz-1);
    REAL C  = 0.00001;
    REAL kx = C/(12*dx*dx);
    REAL ky = C/(12*dy*dy);
    REAL kz = C/(12*dz*dz);
    d_kx = kx; d_ky = ky; d_kz = kz;

    size_t n = (size_t)Nx*Ny*Nz;
    REAL* d_u  = (REAL*)malloc(sizeof(REAL)*n);
    REAL* d_uo = (REAL*)malloc(sizeof(REAL)*n);
    REAL* d_Lu = (REAL*)calloc(n, sizeof(REAL));   /* cudaMemset once (main.cpp:124) */
    memcpy(d_u, u_in, sizeof(REAL)*n);

    dim3 threadsPerBlock3D(blockX,blockY,blockZ);
    unsigned blocksInX = DIVIDE_INTO(Nx,blockX);
    unsigned blocksInY = DIVIDE_INTO(Ny,blockY);
    unsigned blocksInZ = DIVIDE_INTO(Nz,blockZ);
    dim3 numBlocks3D(blocksInX,blocksInY,blocksInZ);
    dim3 threadsPerBlockXY(blockX,blockY);  dim3 numBlocksXY(blocksInX,blocksInY);
    dim3 threadsPerBlockXZ(blockX,blockZ);  dim3 numBlocksXZ(blocksInY,blocksInZ);
    dim3 threadsPerBlockYZ(blockY,blockZ);  dim3 numBlocksYZ(blocksInX,blocksInZ);
    dim3 threadsPerBlock2D(64,4,1);
    dim3 numBlocks2D(DIVIDE_INTO(Nx,64), DIVIDE_INTO(Ny,4), DIVIDE_INTO(Nz,LOOP));

    REAL dt = 0, dt1 = 0, dt2 = 0, it = 0, t = 0;
    while (t < tEnd) {
        dt1 = CFL*dx/1.0;
        dt2 = 1./(2*C*(1/dx/dx+1/dy/dy+1/dz/dz))*0.9;
        dt = dt1 < dt2 ? dt1 : dt2;
        if ((t+dt) > tEnd) { dt = tEnd-t; }
        t += dt; it += 1;
        memcpy(d_uo, d_u, sizeof(REAL)*n);
        for (unsigned step = 1; step <= 3; step++) {
            LAUNCH(numBlocksYZ, threadsPerBlockYZ, Compute_dF(d_u, d_Lu, Nx, Ny, Nz, dx));
            LAUNCH(numBlocksXZ, threadsPerBlockXZ, Compute_dG(d_u, d_Lu, Nx, Ny, Nz, dy));
            LAUNCH(numBlocksXY, threadsPerBlockXY, Compute_dH(d_u, d_Lu, Nx, Ny, Nz, dz));
            if (laplace_mode == 0) {
                LAUNCH(numBlocks2D, threadsPerBlock2D,
                       Compute_Laplace_Async(d_u, d_Lu, Nx, Nx, Ny, Nz, 3, Nz-2, LOOP));
            } else {
                dim3 g(DIVIDE_INTO(Nx,64), DIVIDE_INTO(Ny,4), 1);
                LAUNCH(g, threadsPerBlock2D, Compute_Laplace(d_u, d_Lu, Nx, Ny, Nz));
            }
            LAUNCH(numBlocks3D, threadsPerBlock3D, Compute_RK(d_u, d_uo, d_Lu, step, Nx, Ny, Nz, dt));
        }
    }
    memcpy(u_out, d_u, sizeof(REAL)*n);
    if (dt_out) *dt_out = dt;
    if (it_out) *it_out = it;
    if (t_out)  *t_out  = t;
    free(d_u); free(d_uo); free(d_Lu);
    return 0;
}
