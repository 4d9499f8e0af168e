/* diffusion3d.run -- drop-in for SingleGPU/Diffusion3d_baselineCode (main.cpp:27-152) and, built with
 * -DBLOCKING, for SingleGPU/Diffusion3d_Blocking (8 arguments, no block sizes).
 *   ./diffusion3d.run C L W H Nx Ny Nz iters block_x block_y block_z
 * L, W, H are parsed with atoi exactly like the reference. */
#include <math.h>
#include "common.h"

int main(int argc, char** argv)
{
    double C;
    unsigned L, W, H, Nx, Ny, Nz, max_iters;
#ifdef BLOCKING
    if (argc == 9) {
#else
    if (argc == 12) {
#endif
        C = atof(argv[1]); L = atoi(argv[2]); W = atoi(argv[3]); H = atoi(argv[4]);
        Nx = atoi(argv[5]); Ny = atoi(argv[6]); Nz = atoi(argv[7]); max_iters = atoi(argv[8]);
    } else {
#ifdef BLOCKING
        printf("Usage: %s diffCoef L W H NX NY NZ n_iter\n", argv[0]);
#else
        printf("Usage: %s diffCoef L W H NX NY NZ n_iter block_x block_y block_z\n", argv[0]);
#endif
        exit(1);
    }
    adv_desc d;
    memset(&d, 0, sizeof(d));
#ifdef BLOCKING
    d.target = ADV_TARGET_HEAT3D_SGPU_BLOCKING;
#else
    d.target = ADV_TARGET_HEAT3D_SGPU_BASELINE;
#endif
    d.precision = ADV_F64; d.nx = Nx; d.ny = Ny; d.nz = Nz; d.lx = L; d.ly = W; d.lz = H; d.coef = C; d.nranks = 1;
    adv_solver* s;
    die_on(adv_create(&d, &s), "adv_create");
    const double dx = (double)L / (Nx - 1), dy = (double)W / (Ny - 1), dz = (double)H / (Nz - 1), dt = adv_dt(s);
    const double tFinal = 0.1 + dt * max_iters;
    printf("dx: %g, dy: %g, dz: %g, dt: %g, final time: %g\n\n", dx, dy, dz, dt, tFinal);

    const size_t n = adv_local_elems(s);
    double* h_u = (double*)malloc(sizeof(double) * n);
    die_on(adv_init_reference_ic(s), "adv_init_reference_ic");
    double t0 = now_s();
    die_on(adv_step(s, (int)max_iters), "adv_step");
    die_on(adv_sync(s), "adv_sync");
    double tCPU = now_s() - t0;
    printf("Computation took %lf seconds\n", tCPU);
    die_on(adv_download(s, h_u), "adv_download");

    float gflops = calc_gflops_steps((float)tCPU, max_iters, Nx, Ny, Nz);
    printf("=========================== %s =======================\n", "HeatEq3D (13-pt)");
    printf("Optimization                                 :  %s\n", "fused TMA stage kernel (B200)");
    printf("Kernel time ex. data transfers               :  %lf seconds\n", tCPU);
    printf("===================================================================\n");
    printf("Total effective GFLOPs                       :  %lf\n", gflops);
    printf("===================================================================\n");
    printf("3D Grid Size                                 :  %d x %d x %d \n", Nx, Ny, Nz);
    printf("Iterations                                   :  %d\n", max_iters);
    printf("Final Time                                   :  %g\n", tFinal);
    printf("===================================================================\n");
    /* error norms against the analytic Gaussian, indexed correctly (the reference's CalcError reads u[i]) */
    double l1 = 0, l2 = 0, linf = 0;
    for (unsigned k = 0; k < Nz; k++) for (unsigned j = 0; j < Ny; j++) for (unsigned i = 0; i < Nx; i++) {
        const double x = -5 + i * dx, y = -5 + j * dy, z = -5 + k * dz;
        const double ex = sqrt(0.1 / tFinal) * (0.1 / tFinal) * exp(-(x * x + y * y + z * z) / (4 * tFinal));
        const double err = ex - h_u[i + (size_t)Nx * j + (size_t)Nx * Ny * k];
        l1 += fabs(err); l2 += err * err; linf = fmax(linf, fabs(err));
    }
    printf("L1 norm                                       :  %e\n", dx * dy * dz * l1);
    printf("L2 norm                                       :  %e\n", sqrt(dx * dy * dz * l2));
    printf("Linf norm                                     :  %e\n", linf);
    print_b200_lines(tCPU, max_iters, (double)Nx * Ny * Nz, 8);
    save_binary_f32(h_u, n, "result.bin");
    adv_destroy(s);
    free(h_u);
    return 0;
}
