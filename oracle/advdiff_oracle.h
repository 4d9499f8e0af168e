/* TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement ("oracle") of the reference's explicit time-stepping path:
 * 2-D/3-D heat (4th-order Laplacian) and 3-D Burgers (WENO5 + Lax-Friedrichs
 * split) inside SSP-RK3, with each parity target's frame / coefficient /
 * boundary quirks.  Each routine cites the reference file:line it follows.
 *
 * PARITY PINNED: tests/test_oracle_pin.py checks every routine bit-for-bit
 * against oracle/_ref (the reference's own kernel bodies replayed on the CPU)
 * where /root/reference is present, and against the committed outputs of that
 * replay in tests/golden/ everywhere else.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library, as the checker or the timed
 * CPU baseline.  The product (multigpu_advectiondiffusion_b200, libadvdiff.so)
 * never links, imports or calls it.
 *
 * Suffix _f64 / _f32 selects REAL (the reference's USE_FLOAT switch).
 */
#ifndef ADVDIFF_ORACLE_H
#define ADVDIFF_ORACLE_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ORC_DECL(REAL, S) \
void orc_burgers3d_mgpu_rhs_##S(const REAL* u, REAL* Lu, int nx, int ny, int NZ, REAL dx, REAL dy, REAL dz); \
int  orc_burgers3d_mgpu_step_##S(REAL* u, REAL* uo, REAL* Lu, int nx, int ny, int NZ, REAL dx, REAL dy, REAL dz, REAL dt); \
int  orc_burgers3d_mgpu_run_##S(REAL* u, int nx, int ny, int nz_phys, REAL L, REAL W, REAL H, REAL CFL, REAL tEnd, int max_steps, double* dt_out, double* it_out, double* t_out); \
int  orc_burgers3d_mgpu_run_cfl_##S(REAL* u, int nx, int ny, int nz_phys, REAL L, REAL W, REAL H, REAL CFL, REAL tEnd, int max_steps, double* dt_out, double* it_out, double* t_out); \
int  orc_burgers3d_sgpu_step_##S(REAL* u, REAL* uo, REAL* Lu, int nx, int ny, int nz, REAL dx, REAL dy, REAL dz, REAL kx, REAL ky, REAL kz, REAL dt, int visc_quirk); \
int  orc_burgers3d_sgpu_run_##S(REAL* u, int nx, int ny, int nz, REAL L, REAL W, REAL H, REAL CFL, REAL tEnd, int visc_quirk, int max_steps, double* dt_out, double* it_out, double* t_out); \
int  orc_heat3d_step_##S(REAL* u, REAL* uo, REAL* Lu, int nx, int ny, int nz, REAL kx, REAL ky, REAL kz, REAL dt, int variant); \
int  orc_heat3d_run_##S(REAL* u, int nx, int ny, int nz, REAL L, REAL W, REAL H, REAL C, int iters, int variant, double* dt_out, int* it_out); \
int  orc_heat2d_run_##S(REAL* u, int nx, int ny, REAL L, REAL W, REAL C, int iters, double* dt_out); \
void orc_ic_heat3d_sgpu_##S(REAL* u0, int nx, int ny, int nz, REAL dx, REAL dy, REAL dz); \
void orc_ic_heat2d_sgpu_##S(REAL* u0, int nx, int ny, REAL dx, REAL dy); \
void orc_ic_burgers3d_sgpu_##S(REAL* u0, int nx, int ny, int nz, REAL dx, REAL dy, REAL dz); \
void orc_ic_burgers3d_mgpu_##S(REAL* u0, int nx, int ny, int nz, REAL dx, REAL dy, REAL dz); \
void orc_ic_heat3d_mgpu_##S(REAL* u0, int nx, int ny, int nz); \
int  orc_heat2d_mgpu_run_##S(REAL* u, int nx, int ny_phys, REAL L, REAL W, REAL K, int iters, double* dt_out, int* it_out, double* t_out); \
void orc_burgers2d_mgpu_rhs_##S(const REAL* u, REAL* Lu, int nx, int NY, REAL dx, REAL dy); \
int  orc_burgers2d_mgpu_run_##S(REAL* u, int nx, int ny_phys, REAL L, REAL W, REAL CFL, REAL tEnd, int max_steps, double* dt_out, double* it_out, double* t_out); \
void orc_ic_heat2d_mgpu_##S(REAL* u0, int nx, int ny); \
void orc_ic_burgers2d_mgpu_##S(REAL* u0, int nx, int ny, REAL dx, REAL dy); \
void orc_slab_split_##S(const REAL* h_q, REAL* h_s_q, int n, int nx, int ny, int _Nz); \
void orc_slab_merge_##S(const REAL* h_s_q, REAL* h_q, int n, int nx, int ny, int _Nz); \
int  orc_save_binary_##S(const REAL* u, size_t n, const char* name);

ORC_DECL(double, f64)
ORC_DECL(float, f32)

int orc_num_threads(void);
void orc_set_num_threads(int n);

#ifdef __cplusplus
}
#endif
#endif
