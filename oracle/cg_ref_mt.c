/*
 * cg_ref_mt.c — TEST / BENCH INFRASTRUCTURE ONLY (never linked into the product library).
 *
 * The reference's CPU conjugate gradient (Mgcg/cuBlas/Mgcg/ConjugateGradientCpu.cs:45-98) with the row loops the
 * reference itself marks as parallelisable — its `Parallel.For` lines are present but commented out
 * (SparseMatrix.cs:71, LongVector.cs:21,44) — run over OpenMP threads, so that the timed CPU baseline can use every
 * host core.  Same operation order per row as cg_ref.c (storage-order row sums, -ffp-contract=off); dots are summed as
 * per-thread partials over static contiguous row blocks, combined in thread order (deterministic for a thread count).
 * Used by bench.py's cpu_baseline / --impl reference legs; parity of this file against cg_ref.c is in tests/test_oracle.py.
 */
#include <math.h>
#include <omp.h>
#include <stdlib.h>
#include <string.h>

#define MT_MAX_THREADS 1024

static void mt_spmv(int rows, const double *el, const int *ro, const int *ci, const double *v, double *ans)
{
#pragma omp parallel for schedule(static)
    for (int i = 0; i < rows; i++) {                 /* SparseMatrix.cs:71-87 */
        double s = 0;
        for (int k = ro[i]; k < ro[i + 1]; k++) s += el[k] * v[ci[k]];
        ans[i] = s;
    }
}

static double mt_dot(int n, const double *a, const double *b)
{
    double part[MT_MAX_THREADS];
    int nt = 1;
#pragma omp parallel
    {
        int t = omp_get_thread_num();
#pragma omp single
        nt = omp_get_num_threads();
        double s = 0;
#pragma omp for schedule(static)
        for (int i = 0; i < n; i++) s += a[i] * b[i];   /* LongVector.cs:23-28 */
        part[t] = s;
    }
    double s = 0;
    for (int t = 0; t < nt; t++) s += part[t];
    return s;
}

static void mt_set_added(int n, double *ans, const double *l, const double *r, double a)
{
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n; i++) ans[i] = l[i] + a * r[i];   /* LongVector.cs:46-49 */
}

int ref_mt_max_threads(void) { return omp_get_num_procs(); }

/* Same contract as ref_cg_solve (cg_ref.c) for norm ABS_L2 (0) and REL_L2 (2), non-strict min-iteration.
 * returns 0 = converged, 3 = iteration > max_iteration (the C# throw).  threads <= 0: all cores. */
int ref_cg_solve_mt(int threads, int n, const double *el, const int *ro, const int *ci,
                    double *x, const double *b, double *r, double *p, double *Ap,
                    double allowable, int min_iteration, int max_iteration, int norm, int *iteration, double *residual)
{
    if (threads <= 0) threads = omp_get_num_procs();
    omp_set_num_threads(threads > MT_MAX_THREADS ? MT_MAX_THREADS : threads);
    mt_spmv(n, el, ro, ci, x, Ap);
    mt_set_added(n, r, b, Ap, -1);
    memcpy(p, r, sizeof(double) * (size_t)n);
    double rr = mt_dot(n, r, r);
    const double rr0 = rr;
    int it;
    for (it = 0;; it++) {
        mt_spmv(n, el, ro, ci, p, Ap);
        double alpha = rr / mt_dot(n, p, Ap);
        mt_set_added(n, x, x, p, alpha);
        mt_set_added(n, r, r, Ap, -alpha);
        double rrNew = mt_dot(n, r, r);
        int below;
        if (norm == 2) { *residual = sqrt(rrNew / rr0); below = rrNew / rr0 < allowable * allowable; }
        else { *residual = sqrt(rrNew); below = *residual < allowable; }
        if (it >= min_iteration) {                       /* ConjugateGradient.cs:64-77 */
            if (it > max_iteration) { *iteration = it; return 3; }
            if (below) break;
        }
        double beta = rrNew / rr;
        mt_set_added(n, p, r, p, beta);
        rr = rrNew;
    }
    *iteration = it;
    return 0;
}
