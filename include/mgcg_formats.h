/*
 * mgcg_formats.h — the two other input surfaces of the reference, in front of the same solver.
 *
 *  (1) ELL, the matrix container of the reference's hand-written OpenCL variant
 *      (Mgcg/HandmadeCL/MgcgCL/SparseMatrix.cs:13-47): three arrays
 *          Elements[i*MaxNonzeroCountPerRow + k], ColumnIndeces[i*MaxNonzeroCountPerRow + k], NonzeroCounts[i]
 *      with slot k = 0 of every row holding the DIAGONAL (SparseMatrix.cs:74-92; NonzeroCounts[i] >= 1 always, :52-63)
 *      and the other entries in insertion order (:147-153).  Its solver uploads the three arrays, x and b, iterates with the
 *      max|r_i| stop rule and reads x back (ConjugateGradientSingleGpu.cs:213-296).
 *      Here the ELL arrays are converted ON THE DEVICE to the CSR the kernels stream (slot order kept, so a row's sum runs in
 *      the order of SparseMatrix.Multiply, SparseMatrix.cs:183-205) and the usual solve follows.
 *
 *  (2) the four-call object of the ViennaCL variant (Mgcg/ViennaCL/Mgcg/ComputerGpu.hpp:68-103):
 *          ComputerGpu(n); Write(elements, rowOffsets, columnIndeces, x, b); Solve(residual, minIteration, maxIteration);
 *          Read(x); Iteration()
 *      with UNSIGNED CSR indices (ComputerGpu.hpp:47), the relative stop rule r.r / r0.r0 < residual^2 and the strict
 *      `minIteration < iteration` (ComputerGpu.cpp:66); Iteration() is the number of iterations performed (:51,66).
 *
 * Plain pointers and sizes only.  Status codes and mgcg_last_error as in mgcg.h.
 */
#ifndef MGCG_FORMATS_H
#define MGCG_FORMATS_H

#include "mgcg.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- (1) ELL ------------------------------------------------------------------------------------------------------ */

/* Sum of the row counts = entries of the CSR form; < 0 on error (a count outside [1, max_nonzero_per_row], or a total that
 * does not fit int32).  d_nonzero_counts: device array of `count` ints. */
long long mgcg_ell_nnz(mgcg_ctx *c, int count, int max_nonzero_per_row, const int *d_nonzero_counts);

/* Device ELL -> device CSR (arrays sized by the caller: elements / column_indeces mgcg_ell_nnz entries, row_offsets count + 1).
 * Slot order is kept; unused slots are never read. */
int mgcg_ell_to_csr(mgcg_ctx *c, int count, int max_nonzero_per_row,
                    const double *d_ell_elements, const int *d_ell_column_indeces, const int *d_nonzero_counts,
                    double *d_elements, int *d_row_offsets, int *d_column_indeces);

/* Host ELL arrays in, solution out: upload + conversion + setup + solve + download — the whole of
 * ConjugateGradientSingleGpu.Initialize + Solve of the OpenCL variant (ConjugateGradientSingleGpu.cs:213-296).
 * x holds the initial guess on entry (its bufferX is written from this.x, :218).  For that variant's stop rule pass
 * opt->norm = MGCG_NORM_ABS_LINF. */
int mgcg_solve_host_ell(mgcg_ctx *c, int count, int max_nonzero_per_row,
                        const double *elements, const int *column_indeces, const int *nonzero_counts,
                        const double *b, double *x, const mgcg_options *opt, mgcg_result *res);

/* ---- (2) Write / Solve / Read / Iteration ------------------------------------------------------------------------- */

typedef struct mgcg_computer mgcg_computer;

/* ComputerGpu(int n) (ComputerGpu.cpp:8-21): a solver object for n unknowns on `device`.  NULL on failure. */
mgcg_computer *mgcg_computer_create(int device, int n);
void           mgcg_computer_destroy(mgcg_computer *k);

/* Optional, before Solve: precondition with the library's multigrid (nx, ny, nz: structured-grid hint or 0).  The reference
 * object has no preconditioner (its Jacobi-preconditioned ViennaCL call is commented out, ComputerGpu.cpp:86-91). */
int mgcg_computer_set_preconditioner(mgcg_computer *k, int precond, int nx, int ny, int nz);

/* Write (ComputerGpu.cpp:23-35, ComputerGpu.hpp:84): host CSR with unsigned indices, initial x and right-hand side. */
int mgcg_computer_write(mgcg_computer *k, const double *elements, const unsigned int *row_offsets,
                        const unsigned int *column_indeces, int elements_count, const double *x, const double *b);

/* Solve (ComputerGpu.cpp:37-95): CG from the written x until  minIteration < iteration  and  r.r / r0.r0 < residual^2.
 * The reference accepts maxIteration and never looks at it; here a positive value bounds the loop
 * (MGCG_ERR_NOT_CONVERGED), zero or negative means unbounded like the reference. */
int mgcg_computer_solve(mgcg_computer *k, double residual, int min_iteration, int max_iteration);

/* Read (ComputerGpu.cpp:97-104): the solution to a host array of n doubles. */
int mgcg_computer_read(mgcg_computer *k, double *x);

/* Iteration (ComputerGpu.hpp:62): iterations performed by the last Solve. */
int mgcg_computer_iteration(const mgcg_computer *k);

/* the last Solve's full record (times, residual, levels) */
int mgcg_computer_result(const mgcg_computer *k, mgcg_result *res);

#ifdef __cplusplus
}
#endif
#endif
