/* diffusion2d.run -- drop-in for SingleGPU/Diffusion2d (main.cpp:27-155).
 *   ./diffusion2d.run C L W Nx Ny iters block_x block_y */
#include <math.h>
#include "common.h"

int main(int argc, char** argv)
{
    double C;
    unsigned L, W, Nx, Ny, max_iters;
    if (argc == 9) {
        C = atof(argv[1]); L = atoi(argv[2]); W = atoi(argv[3]); Nx = atoi(argv[4]); Ny = atoi(argv[5]); max_iters = atoi(argv[6]);
    } else {
        printf("Usage: %s PO CFL L W NX NY n_iter block_x block_y\n", argv[0]);
        exit(1);
    }
    adv_desc d;
    memset(&d, 0, sizeof(d));
    d.target = ADV_TARGET_HEAT2D_SGPU; d.precision = ADV_F64; d.nx = Nx; d.ny = Ny; d.nz = 1; d.lx = L; d.ly = W; d.coef = C; d.nranks = 1;
    adv_solver* s;
    die_on(adv_create(&d, &s), "adv_create");
    const double dx = (double)L / (Nx - 1), dy = (double)W / (Ny - 1), dt = adv_dt(s), tFinal = 0.1 + dt * max_iters;
    printf("dx: %g, dy: %g, dt: %g, final time: %g\n\n", dx, dy, dt, tFinal);
    const size_t n = adv_local_elems(s);
    double* h_u = (double*)malloc(sizeof(double) * n);
    die_on(adv_init_reference_ic(s), "adv_init_reference_ic");
    double t0 = now_s();
    die_on(adv_step(s, (int)max_iters), "adv_step");
    die_on(adv_sync(s), "adv_sync");
    double tCPU = now_s() - t0;
    printf("Computation took %lf seconds\n", tCPU);
    die_on(adv_download(s, h_u), "adv_download");
    float gflops = max_iters * (double)((Nx * Ny) * 1e-9 * FLOPS) / tCPU;
    printf("=========================== %s =======================\n", "HeatEq2D (13-pt)");
    printf("Optimization                                 :  %s\n", "fused stage kernel (B200)");
    printf("Kernel time ex. data transfers               :  %lf seconds\n", tCPU);
    printf("===================================================================\n");
    printf("Total effective GFLOPs                       :  %lf\n", gflops);
    printf("===================================================================\n");
    printf("2D Grid Size                                 :  %d x %d \n", Nx, Ny);
    printf("Iterations                                   :  %d\n", max_iters);
    printf("Final Time                                   :  %g\n", tFinal);
    printf("===================================================================\n");
    print_b200_lines(tCPU, max_iters, (double)Nx * Ny, 8);
    save_binary_f32(h_u, n, "result.bin");
    adv_destroy(s);
    free(h_u);
    return 0;
}
