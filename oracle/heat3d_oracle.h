/* heat3d_oracle — CPU restatement of the reference's heat3D hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this; libheat3d.so and the heat3D
 * driver never do (they fail loudly without a GPU instead of falling back to this).
 *
 * Parity status: the reference ships no tests, golden vectors or valid outputs for this
 * path (its only artefact, output/out.dat, is an all-zero field) — "parity unpinned" by
 * the reference's own fixtures.  This restatement is instead pinned, bit for bit, against
 * the reference's own host program compiled here (oracle/_ref/heat3D_ref, see
 * oracle/Makefile + ref_patch.sed): tests/test_oracle.py and tests/golden/.
 *
 * All citations are into /root/reference/HeatEquation3D/src/ ("cu" = heat3D.cu,
 * "cpp" = heat3D.cpp).  Storage here is one flat array, x fastest:
 *     idx(i,j,k) = (k*ny + j)*nx + i        (i=x, j=y, k=z)
 * which is the order the reference's dump walks the field (cpp:1065-1067); the
 * reference's own T[i][j][k] has the same (x,y,z) meaning.
 */
#ifndef HEAT3D_ORACLE_H
#define HEAT3D_ORACLE_H
#ifdef __cplusplus
extern "C" {
#endif

enum { H3O_STRICT = 0, H3O_FMA = 1 };

/* spacing, dt, Dx/Dy/Dz exactly as cu:352-367 (alpha = 1.0 cu:345, CFL = 0.4 cu:348, L = 1 cu:339-341). */
void h3o_coefficients(unsigned nx, unsigned ny, unsigned nz, double sp[3], double* dt, double D[3]);

/* IC + Dirichlet shell, in the reference's write order (cu:409-453): zero; y-top plane = 1.0;
 * then x=0, x=nx-1, z=0, z=nz-1 faces = j*sp_y (overwriting the top edges). */
void h3o_init_field(double* T, unsigned nx, unsigned ny, unsigned nz, double sp_y);

/* One FTCS update of planes z in [z0,z1) ∩ [1,nz-2] (interior cells only); every other cell of
 * those planes is copied src->dst.  Expression cu:128-131.  arith = H3O_STRICT: host evaluation
 * order, one rounding per operation; H3O_FMA: the contraction nvcc emits for the same expression. */
void h3o_step_planes(const double* src, double* dst, unsigned nx, unsigned ny, unsigned nz,
                     double Dx, double Dy, double Dz, unsigned z0, unsigned z1, int arith, int threads);

/* nsteps full-domain steps in place (scratch: same size as T). */
void h3o_steps(double* T, double* scratch, unsigned nx, unsigned ny, unsigned nz,
               double Dx, double Dy, double Dz, unsigned nsteps, int arith, int threads);

/* max(DBL_MIN, max over interior |T - T0|)  — cpp:915-920. */
double h3o_residual(const double* T, const double* T0, unsigned nx, unsigned ny, unsigned nz);

/* The reference's time loop (cpp:497-970): returns number of updates applied; *final_iters and
 * *converged as the reference reports them (cpp:966-969). */
unsigned h3o_run(double* T, double* scratch, unsigned nx, unsigned ny, unsigned nz,
                 double Dx, double Dy, double Dz, unsigned iter_max, double eps,
                 unsigned* final_iters, int* converged, int arith, int threads);

/* mean over interior of sqrt(pow(T - j*sp_y, 2)), summed k-outer, j, i-inner — cpp:992-998. */
double h3o_mean_abs_error_vs_y(const double* T, unsigned nx, unsigned ny, unsigned nz, double sp_y);

/* Single-rank Tecplot ASCII dump, cpp:1059-1075.  Returns 0 on success. */
int h3o_write_dump(const char* path, const double* T, unsigned nx, unsigned ny, unsigned nz, const double sp[3]);

#ifdef __cplusplus
}
#endif
#endif
