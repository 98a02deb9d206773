/* See heat3d_oracle.h — TEST INFRASTRUCTURE ONLY; CPU restatement of the reference hot path. */
#include "heat3d_oracle.h"
#include <float.h>
#include <math.h>
#include <stdio.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static double min3(double a, double b, double c) { double m = a < b ? a : b; return m < c ? m : c; }

void h3o_coefficients(unsigned nx, unsigned ny, unsigned nz, double sp[3], double* dt, double D[3])
{
    const double alpha = 1.0, CFL = 0.4;                 /* cu:345, cu:348 */
    const double len[3] = { 1.0, 1.0, 1.0 };             /* cu:339-341 */
    const unsigned n[3] = { nx, ny, nz };
    for (int d = 0; d < 3; ++d) sp[d] = len[d] / (double)(n[d] - 1.0);            /* cu:353-355 */
    *dt = CFL * 1.0 / (3 * 2) * pow(min3(sp[0], sp[1], sp[2]), 2.0) / alpha;      /* cu:359-360 */
    for (int d = 0; d < 3; ++d) D[d] = *dt * alpha / (pow(sp[d], 2.0));           /* cu:365-367 */
}

#define IDX(i, j, k) (((size_t)(k) * ny + (j)) * nx + (i))

void h3o_init_field(double* T, unsigned nx, unsigned ny, unsigned nz, double sp_y)
{
    memset(T, 0, (size_t)nx * ny * nz * sizeof(double));                          /* cu:409-412 */
    for (unsigned k = 0; k < nz; ++k)                                             /* cu:416-420 */
        for (unsigned i = 0; i < nx; ++i) T[IDX(i, ny - 1, k)] = 1.0;
    for (unsigned k = 0; k < nz; ++k)                                             /* cu:425-437 */
        for (unsigned j = 0; j < ny; ++j) {
            T[IDX(0, j, k)] = j * sp_y;
            T[IDX(nx - 1, j, k)] = j * sp_y;
        }
    for (unsigned j = 0; j < ny; ++j)                                             /* cu:441-453 */
        for (unsigned i = 0; i < nx; ++i) {
            T[IDX(i, j, 0)] = j * sp_y;
            T[IDX(i, j, nz - 1)] = j * sp_y;
        }
}

/* cu:128-131 as the host compiler evaluates it: (a - 2.0*c) + b per axis, then
 * ((c + Dx*sx) + Dy*sy) + Dz*sz; every operation rounded (build has -ffp-contract=off). */
static inline double update_strict(double c, double xm, double xp, double ym, double yp, double zm, double zp,
                                   double Dx, double Dy, double Dz)
{
    return c + Dx * (xp - 2.0 * c + xm) + Dy * (yp - 2.0 * c + ym) + Dz * (zp - 2.0 * c + zm);
}

/* The same expression as nvcc contracts it (PTX of computeT embedded in the reference's
 * heat3D_generated_heat3D.cu.o): c2 = c+c; r = fma((p - c2) + m, D, r) chained x, y, z. */
static inline double update_fma(double c, double xm, double xp, double ym, double yp, double zm, double zp,
                                double Dx, double Dy, double Dz)
{
    const double c2 = c + c;
    double r = fma((xp - c2) + xm, Dx, c);
    r = fma((yp - c2) + ym, Dy, r);
    r = fma((zp - c2) + zm, Dz, r);
    return r;
}

void h3o_step_planes(const double* src, double* dst, unsigned nx, unsigned ny, unsigned nz,
                     double Dx, double Dy, double Dz, unsigned z0, unsigned z1, int arith, int threads)
{
    if (z1 > nz) z1 = nz;
    (void)threads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(threads > 0 ? threads : 1)
#endif
    for (long kk = (long)z0; kk < (long)z1; ++kk) {
        const unsigned k = (unsigned)kk;
        const double* s = src + IDX(0, 0, k);
        double* d = dst + IDX(0, 0, k);
        if (k == 0 || k == nz - 1) { memcpy(d, s, (size_t)nx * ny * sizeof(double)); continue; }
        const size_t sy = nx, sz = (size_t)nx * ny;
        for (unsigned j = 0; j < ny; ++j) {
            const double* r = s + (size_t)j * nx;
            double* o = d + (size_t)j * nx;
            if (j == 0 || j == ny - 1) { memcpy(o, r, (size_t)nx * sizeof(double)); continue; }
            o[0] = r[0];
            o[nx - 1] = r[nx - 1];
            if (arith == H3O_FMA)
                for (unsigned i = 1; i + 1 < nx; ++i)
                    o[i] = update_fma(r[i], r[i - 1], r[i + 1], r[i - sy], r[i + sy], r[i - sz], r[i + sz], Dx, Dy, Dz);
            else
                for (unsigned i = 1; i + 1 < nx; ++i)
                    o[i] = update_strict(r[i], r[i - 1], r[i + 1], r[i - sy], r[i + sy], r[i - sz], r[i + sz], Dx, Dy, Dz);
        }
    }
}

void h3o_steps(double* T, double* scratch, unsigned nx, unsigned ny, unsigned nz,
               double Dx, double Dy, double Dz, unsigned nsteps, int arith, int threads)
{
    double* a = T; double* b = scratch;
    for (unsigned s = 0; s < nsteps; ++s) {
        h3o_step_planes(a, b, nx, ny, nz, Dx, Dy, Dz, 0, nz, arith, threads);
        double* t = a; a = b; b = t;
    }
    if (a != T) memcpy(T, a, (size_t)nx * ny * nz * sizeof(double));
}

double h3o_residual(const double* T, const double* T0, unsigned nx, unsigned ny, unsigned nz)
{
    double res = DBL_MIN;                                                         /* cpp:915 */
    for (unsigned k = 1; k + 1 < nz; ++k)
        for (unsigned j = 1; j + 1 < ny; ++j)
            for (unsigned i = 1; i + 1 < nx; ++i) {
                const double dlt = fabs(T[IDX(i, j, k)] - T0[IDX(i, j, k)]);      /* cpp:919-920 */
                if (dlt > res) res = dlt;
            }
    return res;
}

unsigned h3o_run(double* T, double* scratch, unsigned nx, unsigned ny, unsigned nz,
                 double Dx, double Dy, double Dz, unsigned iter_max, double eps,
                 unsigned* final_iters, int* converged, int arith, int threads)
{
    double norm = 1.0;                                                            /* cpp:281 */
    double* cur = T; double* prev = scratch;
    unsigned applied = 0;
    *final_iters = 0; *converged = 0;
    for (unsigned time = 0; time < iter_max; ++time) {                            /* cpp:497 */
        double* t = cur; cur = prev; prev = t;          /* rotate instead of the T0=T copy, cpp:502-505 */
        h3o_step_planes(prev, cur, nx, ny, nz, Dx, Dy, Dz, 0, nz, arith, threads);
        ++applied;
        const double res = h3o_residual(cur, prev, nx, ny, nz);
        if (time == 0 && res != 0.0) norm = res;                                  /* cpp:926-928 */
        if (res / norm < eps) { *converged = 1; *final_iters = time; break; }     /* cpp:951-969 */
    }
    if (cur != T) memcpy(T, cur, (size_t)nx * ny * nz * sizeof(double));
    return applied;
}

double h3o_mean_abs_error_vs_y(const double* T, unsigned nx, unsigned ny, unsigned nz, double sp_y)
{
    double error = 0.0;
    for (unsigned k = 1; k + 1 < nz; ++k)                                         /* cpp:994-997 */
        for (unsigned j = 1; j + 1 < ny; ++j)
            for (unsigned i = 1; i + 1 < nx; ++i)
                error += sqrt(pow(T[IDX(i, j, k)] - j * sp_y, 2.0));
    error /= ((nx - 2) * (ny - 2) * (nz - 2));            /* unsigned product, as cpp:998 */
    return error;
}

int h3o_write_dump(const char* path, const double* T, unsigned nx, unsigned ny, unsigned nz, const double sp[3])
{
    FILE* f = fopen(path, "w");
    if (!f) return -1;
    fprintf(f, "TITLE=\"out\"\n");                                                /* cpp:1062-1064 */
    fprintf(f, "VARIABLES = \"X\", \"Y\", \"Z\", \"T\"\n");
    fprintf(f, "ZONE T = \"0\", I=%u, J=%u, K=%u, F=POINT\n", nx, ny, nz);
    for (unsigned k = 0; k < nz; ++k)                                             /* cpp:1065-1073 */
        for (unsigned j = 0; j < ny; ++j)
            for (unsigned i = 0; i < nx; ++i)
                fprintf(f, "%15.5e%15.5e%15.5e%15.5e\n", i * sp[0], j * sp[1], k * sp[2], T[IDX(i, j, k)]);
    return fclose(f);
}
