/* burgers3d.run -- drop-in for SingleGPU/Burgers3d_WENO5 (main.cpp:27-182): same argv, stdout, result.bin.
 *   ./burgers3d.run tEnd CFL L W H Nx Ny Nz block_x block_y block_z
 * block_* are accepted and ignored (tiling is fixed by the fused kernel).  WRITE defaults to 0 like the
 * reference (Burgers.h:32); set ADV_WRITE=1 to get result.bin.  ADV_VISC_ONCE=1 adds the viscous term once
 * per plane instead of reproducing the reference's double count on planes 16..18. */
#include "common.h"

int main(int argc, char** argv)
{
    double tEnd, CFL, L, W, H;
    unsigned Nx, Ny, Nz;
    if (argc == 12) {
        tEnd = atof(argv[1]); CFL = atof(argv[2]); L = atof(argv[3]); W = atof(argv[4]); H = atof(argv[5]);
        Nx = atoi(argv[6]); Ny = atoi(argv[7]); Nz = atoi(argv[8]);
    } else {
        printf("Usage: %s tEnd CFL L W H NX NY NZ block_x block_y block_z\n", argv[0]);
        exit(1);
    }
    const double dx = L / (Nx - 1), dy = W / (Ny - 1), dz = H / (Nz - 1);
    printf("dx: %g, dy: %g, dz: %g, final time: %g\n\n", dx, dy, dz, tEnd);

    adv_desc d;
    memset(&d, 0, sizeof(d));
    d.target = ADV_TARGET_BURGERS3D_SGPU; d.precision = ADV_F64;
    d.nx = Nx; d.ny = Ny; d.nz = Nz; d.lx = L; d.ly = W; d.lz = H; d.coef = CFL;
    d.flags = getenv("ADV_VISC_ONCE") ? ADV_FLAG_VISC_ONCE : 0;
    d.device = 0; d.nranks = 1;
    adv_solver* s;
    die_on(adv_create(&d, &s), "adv_create");
    const size_t n = adv_local_elems(s);
    double* h_u = (double*)malloc(sizeof(double) * n);

    double t0 = now_s();
    die_on(adv_init_reference_ic(s), "adv_init_reference_ic");      /* Call_Init3d case 3 + H2D */
    double HtD = now_s() - t0;

    double dt = 0, it = 0, t = 0;
    t0 = now_s();
    die_on(adv_run_until(s, tEnd, -1, &dt, &it, &t), "adv_run_until");
    die_on(adv_sync(s), "adv_sync");
    double compute = now_s() - t0;
    printf("Computation took %1.6f seconds\n\n", compute);
    printf("dt: %g, iterations: %g, final time: %g\n\n", dt, it, t);

    t0 = now_s();
    die_on(adv_download(s, h_u), "adv_download");
    double DtH = now_s() - t0;
    if (getenv("ADV_WRITE")) save_binary_f32(h_u, n, "result.bin");

    float gflops = calc_gflops_stages((float)compute, (unsigned)it, Nx, Ny, Nz);
    print_summary_mgpu("===========================", "Burgers-3D GPU-WENO5", "=======================", "fused WENO5 stage kernel (B200)", compute, HtD, DtH, gflops, (int)it, Nx, Ny, Nz);
    print_b200_lines(compute, it, (double)Nx * Ny * Nz, 8);
    adv_destroy(s);
    free(h_u);
    return 0;
}
