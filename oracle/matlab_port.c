/* TEST INFRASTRUCTURE ONLY ("Oracle B").
 *
 * C transcription of the reference's Matlab prototypes -- its CPU implementation of the same path -- kept
 * independent of advdiff_oracle.c (which restates the CUDA programs): same operators, Matlab's own operand
 * order, boundary treatment (edge-replicated ghost cells, zeroed outer layer for heat) and, optionally,
 * its dynamic time step dt = CFL*dx/max|u|.  No Matlab/Octave exists on this image; the heat routines are
 * pinned instead to the one executed Matlab run the reference recorded: they reproduce every printed digit of
 * the L1 / Linf tables in Matlab_Prototipes/DiffusionNd/TestingAccuracy.log:49-72 (tests/test_oracle_b.py).
 * The Burgers routine has no such record ("parity unpinned" for mlb_burgers3d_run).  Use: a second opinion --
 * it must agree with Oracle A to rounding wherever the two boundary treatments cannot have reached yet.
 * Never linked or imported by the product.
 *
 *   heat3d.m:48-77 + Laplace3d.m:32-60 ('OMP' strategy)     -> mlb_heat3d_run
 *   heat2d.m:47-75 + Laplace2d.m:6-16                        -> mlb_heat2d_run
 *   LFWENO5FDM3d.m:63-97 + WENO5resAdv_{X,Y,Z}.m:50-125      -> mlb_burgers3d_run
 * Arrays: x fastest, o = i + nx*j + nx*ny*k (the axis names are immaterial to the arithmetic).
 */
#include <math.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>

#define IDX(i,j,k) ((size_t)(i) + (size_t)nx*(size_t)(j) + (size_t)nx*(size_t)ny*(size_t)(k))

/* Laplace3d.m 'OMP' branch: 1-based i,j in (2, n-1), k = 3..nz-2  ==  0-based [2, n-3] in all three */
static void mlb_laplace3d(const double* u, double* Lu, int nx, int ny, int nz, double Dx, double Dy, double Dz)
{
    const ptrdiff_t xy = (ptrdiff_t)nx*ny;
    memset(Lu, 0, sizeof(double)*(size_t)nx*ny*nz);
    #pragma omp parallel for schedule(static)
    for (int k = 2; k <= nz-3; k++)
        for (int j = 2; j <= ny-3; j++)
            for (int i = 2; i <= nx-3; i++) {
                const size_t o = IDX(i,j,k);
                const double center = u[o];
                Lu[o] = Dx/12*(-  u[o-2]  +16*u[o-1] -30*center+16*u[o+1] -u[o+2]   )+
                        Dy/12*(-u[o-2*nx]+16*u[o-nx]-30*center+16*u[o+nx]-u[o+2*nx])+
                        Dz/12*(-u[o-2*xy]+16*u[o-xy]-30*center+16*u[o+xy]-u[o+2*xy]);
            }
}

/* heat3d.m:48-77: fixed dt, `nsteps` steps (the script's while-loop with the clipped last step is the
 * caller's business), BCs zeroed after every step. */
int mlb_heat3d_run(double* u, int nx, int ny, int nz, double Dx, double Dy, double Dz, double dt, int nsteps)
{
    const size_t n = (size_t)nx*ny*nz;
    double* uo = (double*)malloc(n*sizeof(double));
    double* dF = (double*)malloc(n*sizeof(double));
    if (!uo || !dF) { free(uo); free(dF); return 1; }
    for (int it = 0; it < nsteps; it++) {
        memcpy(uo, u, n*sizeof(double));
        mlb_laplace3d(u, dF, nx, ny, nz, Dx, Dy, Dz);
        #pragma omp parallel for schedule(static)
        for (size_t o = 0; o < n; o++) u[o] = uo[o]+dt*dF[o];
        mlb_laplace3d(u, dF, nx, ny, nz, Dx, Dy, Dz);
        #pragma omp parallel for schedule(static)
        for (size_t o = 0; o < n; o++) u[o] = 0.75*uo[o]+0.25*(u[o]+dt*dF[o]);
        mlb_laplace3d(u, dF, nx, ny, nz, Dx, Dy, Dz);
        #pragma omp parallel for schedule(static)
        for (size_t o = 0; o < n; o++) u[o] = (uo[o]+2*(u[o]+dt*dF[o]))/3;
        #pragma omp parallel for schedule(static)
        for (int k = 0; k < nz; k++) for (int j = 0; j < ny; j++) for (int i = 0; i < nx; i++)
            if (i==0 || i==nx-1 || j==0 || j==ny-1 || k==0 || k==nz-1) u[IDX(i,j,k)] = 0;
    }
    free(uo); free(dF);
    return 0;
}

/* heat2d.m:47-75 + Laplace2d.m */
int mlb_heat2d_run(double* u, int nx, int ny, double Dx, double Dy, double dt, int nsteps)
{
    const size_t n = (size_t)nx*ny;
    double* uo = (double*)malloc(n*sizeof(double));
    double* dF = (double*)calloc(n, sizeof(double));
    if (!uo || !dF) { free(uo); free(dF); return 1; }
    for (int it = 0; it < nsteps; it++) {
        memcpy(uo, u, n*sizeof(double));
        for (int stage = 1; stage <= 3; stage++) {
            for (int j = 0; j < ny; j++) for (int i = 0; i < nx; i++) {
                const size_t o = (size_t)i + (size_t)nx*j;
                if (i>=2 && i<=nx-3 && j>=2 && j<=ny-3)
                    dF[o] = Dx/12*(-u[o+2]+16*u[o+1]-30*u[o]+16*u[o-1]-u[o-2]) +
                            Dy/12*(-u[o+2*nx]+16*u[o+nx]-30*u[o]+16*u[o-nx]-u[o-2*nx]);
                else dF[o] = 0.0;
            }
            for (size_t o = 0; o < n; o++)
                u[o] = stage == 1 ? uo[o]+dt*dF[o] : stage == 2 ? 0.75*uo[o]+0.25*(u[o]+dt*dF[o]) : (uo[o]+2*(u[o]+dt*dF[o]))/3;
        }
        for (int j = 0; j < ny; j++) for (int i = 0; i < nx; i++)
            if (i==0 || i==nx-1 || j==0 || j==ny-1) u[(size_t)i + (size_t)nx*j] = 0;
    }
    free(uo); free(dF);
    return 0;
}

/* WENO5resAdv_X.m:50-125 along one line of N cells with element stride s; res[c*s] = df/dx at cell c.
 * W = [w1 w1 w(1..N) wN wN wN]; hn, hp are zero outside I, so the first cell sees a zero left flux. */
static void mlb_weno5_line(const double* w, double* res, ptrdiff_t s, int N, double h, double* W, double* V, double* U, double* hn, double* hp)
{
    W[0] = W[1] = w[0];
    for (int c = 0; c < N; c++) W[c+2] = w[(ptrdiff_t)c*s];
    W[N+2] = W[N+3] = W[N+4] = w[(ptrdiff_t)(N-1)*s];
    for (int q = 0; q < N+5; q++) {
        const double a = fabs(W[q]);                       /* a = |dflux(W)|, dflux = w */
        const double f = W[q]*W[q]/2;                      /* flux = w.^2/2 */
        V[q] = 0.5*(f+a*W[q]);
        U[q] = 0.5*(f-a*W[q]);
        hn[q] = 0; hp[q] = 0;
    }
    const double d0n = 1./10, d1n = 6./10, d2n = 3./10, d0p = 3./10, d1p = 6./10, d2p = 1./10, epsilon = 1e-6;
    for (int I = 2; I <= N+1; I++) {
        const double vmm = V[I-2], vm = V[I-1], v = V[I], vp = V[I+1], vpp = V[I+2];
        const double B0n = 13./12*pow(vmm-2*vm+v,2) + 1./4*pow(vmm-4*vm+3*v,2);
        const double B1n = 13./12*pow(vm-2*v+vp,2) + 1./4*pow(vm-vp,2);
        const double B2n = 13./12*pow(v-2*vp+vpp,2) + 1./4*pow(3*v-4*vp+vpp,2);
        const double alpha0n = d0n/pow(epsilon+B0n,2), alpha1n = d1n/pow(epsilon+B1n,2), alpha2n = d2n/pow(epsilon+B2n,2);
        const double alphasumn = alpha0n+alpha1n+alpha2n;
        const double w0n = alpha0n/alphasumn, w1n = alpha1n/alphasumn, w2n = alpha2n/alphasumn;
        hn[I] = w0n*(2*vmm-7*vm+11*v)/6 + w1n*(-vm+5*v+2*vp)/6 + w2n*(2*v+5*vp-vpp)/6;
        const double umm = U[I-1], um = U[I], u = U[I+1], up = U[I+2], upp = U[I+3];
        const double B0p = 13./12*pow(umm-2*um+u,2) + 1./4*pow(umm-4*um+3*u,2);
        const double B1p = 13./12*pow(um-2*u+up,2) + 1./4*pow(um-up,2);
        const double B2p = 13./12*pow(u-2*up+upp,2) + 1./4*pow(3*u-4*up+upp,2);
        const double alpha0p = d0p/pow(epsilon+B0p,2), alpha1p = d1p/pow(epsilon+B1p,2), alpha2p = d2p/pow(epsilon+B2p,2);
        const double alphasump = alpha0p+alpha1p+alpha2p;
        const double w0p = alpha0p/alphasump, w1p = alpha1p/alphasump, w2p = alpha2p/alphasump;
        hp[I] = w0p*(-umm+5*um+2*u)/6 + w1p*(2*um+5*u-up)/6 + w2p*(11*u-7*up+2*upp)/6;
    }
    for (int I = 2; I <= N+1; I++)
        res[(ptrdiff_t)(I-2)*s] = (hp[I]+hn[I]-hp[I-1]-hn[I-1])/h;
}

static void mlb_res_axis(const double* u, double* res, int nx, int ny, int nz, int axis, double h)
{
    const int N = axis == 0 ? nx : axis == 1 ? ny : nz;
    const ptrdiff_t s = axis == 0 ? 1 : axis == 1 ? nx : (ptrdiff_t)nx*ny;
    const int na = axis == 0 ? ny : nx, nb = axis == 2 ? ny : nz;     /* the two other extents */
    #pragma omp parallel
    {
        double* buf = (double*)malloc(sizeof(double)*5*(size_t)(N+5));
        #pragma omp for collapse(2) schedule(static)
        for (int b = 0; b < nb; b++)
            for (int a = 0; a < na; a++) {
                size_t o;
                if (axis == 0) o = IDX(0, a, b);
                else if (axis == 1) o = IDX(a, 0, b);
                else o = IDX(a, b, 0);
                mlb_weno5_line(u + o, res + o, s, N, h, buf, buf+(N+5), buf+2*(N+5), buf+3*(N+5), buf+4*(N+5));
            }
        free(buf);
    }
}

/* LFWENO5FDM3d.m:63-97.  dynamic_dt = 1: dt = CFL*dx/max|u| as the script; 0: max|u| := 1 (what the CUDA
 * drivers assume, main.c:193), for comparison with Oracle A.  Stops at tEnd or after max_steps (< 0: none). */
int mlb_burgers3d_run(double* u, int nx, int ny, int nz, double dx, double dy, double dz, double CFL, double tEnd,
                      int dynamic_dt, int max_steps, double* dt_out, double* it_out, double* t_out)
{
    const size_t n = (size_t)nx*ny*nz;
    double *uo = (double*)malloc(n*sizeof(double)), *dF = (double*)malloc(n*sizeof(double)),
           *dG = (double*)malloc(n*sizeof(double)), *dH = (double*)malloc(n*sizeof(double));
    if (!uo || !dF || !dG || !dH) { free(uo); free(dF); free(dG); free(dH); return 1; }
    double t = 0, dt = 0; int it = 0;
    while (t < tEnd && (max_steps < 0 || it < max_steps)) {
        double umax = 1.0;
        if (dynamic_dt) { umax = 0; for (size_t o = 0; o < n; o++) if (fabs(u[o]) > umax) umax = fabs(u[o]); }
        dt = CFL*dx/umax; if (t+dt > tEnd) dt = tEnd-t;
        t = t+dt; it = it+1;
        memcpy(uo, u, n*sizeof(double));
        for (int stage = 1; stage <= 3; stage++) {
            mlb_res_axis(u, dF, nx, ny, nz, 0, dx);
            mlb_res_axis(u, dG, nx, ny, nz, 1, dy);
            mlb_res_axis(u, dH, nx, ny, nz, 2, dz);
            #pragma omp parallel for schedule(static)
            for (size_t o = 0; o < n; o++) {
                const double r = u[o]-dt*(dF[o]+dG[o]+dH[o]);
                u[o] = stage == 1 ? uo[o]-dt*(dF[o]+dG[o]+dH[o]) : stage == 2 ? 0.75*uo[o]+0.25*r : (uo[o]+2*r)/3;
            }
        }
    }
    if (dt_out) *dt_out = dt;
    if (it_out) *it_out = it;
    if (t_out) *t_out = t;
    free(uo); free(dF); free(dG); free(dH);
    return 0;
}
