/*
 * oracle/cg_ref.c — TEST INFRASTRUCTURE ONLY (never linked into, imported by or called from the product path).
 *
 * A plain-C, single-threaded, fp64 restatement of the reference's only complete CPU conjugate-gradient
 * solver and of the integer partition / halo rules of its multi-GPU driver.  Each function cites the
 * reference file:line it follows (paths relative to /root/reference).  Compile with -ffp-contract=off so
 * that `a + b*c` is a rounded multiply followed by a rounded add, like the .NET JIT the reference ran on.
 *
 * PARITY PINNING: the reference holds no golden vectors, tests or fixtures, and none
 * of its implementations can be compiled or run in this image (C#/.NET, R, boost, CUDA-5 cuSPARSE, Windows).
 * This restatement is therefore pinned only by (i) the known-answer cases derivable from the reference's own
 * inputs (tridiagonal [1 2 1] against a direct solve, R/CG.R:34 idea) and (ii) line-by-line correspondence
 * with the cited source; tests/test_oracle.py checks (i).  => "parity unpinned" in the strict sense.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ---- stop rules ------------------------------------------------------------------ */
enum { REF_NORM_ABS_L2 = 0,   /* Mgcg/cuBlas/Mgcg/ConjugateGradientCpu.cs:77, Mgcg.cu:251, SimpleConjugateGradient.cu:106 */
       REF_NORM_ABS_LINF = 1, /* Mgcg/HandmadeCL/MgcgCL/ConjugateGradientCpu.cs:75 */
       REF_NORM_REL_L2 = 2 }; /* Mgcg/ViennaCL/Mgcg/ComputerCpu.cpp:82 */

enum { REF_OK = 0, REF_MAX_ITERATION_EXCEEDED = 3 };

/* SparseMatrix.Multiply — Mgcg/cuBlas/Mgcg/SparseMatrix.cs:68-88.
 * answer[i] = 0; then += Elements[k]*vector[Col[k]] in STORAGE order (the generator stores the diagonal first). */
void ref_spmv(int rows, const double *elements, const int *row_offsets, const int *column_indeces,
              const double *vector, double *answer)
{
    for (int i = 0; i < rows; i++) {
        answer[i] = 0;
        for (int k = row_offsets[i]; k < row_offsets[i + 1]; k++) {
            int j = column_indeces[k];
            answer[i] += elements[k] * vector[j];
        }
    }
}

/* LongVectorExtensions.Dot — Mgcg/cuBlas/Mgcg/LongVector.cs:15-31 (sequential left-to-right). */
double ref_dot(int n, const double *left, const double *right)
{
    double answer = 0;
    for (int i = 0; i < n; i++) answer += left[i] * right[i];
    return answer;
}

/* LongVectorExtensions.SetAdded — LongVector.cs:41-51: answer = left + a*right (answer may alias left). */
void ref_set_added(int n, double *answer, const double *left, const double *right, double a)
{
    for (int i = 0; i < n; i++) answer[i] = left[i] + a * right[i];
}

/* LongVectorExtensions.MaxAbsolute — LongVector.cs:58-72. */
double ref_max_absolute(int n, const double *v)
{
    double max = fabs(v[0]);
    for (int i = 1; i < n; i++) { double a = fabs(v[i]); max = a > max ? a : max; }
    return max;
}

/* ConjugateGradient.IsConverged — Mgcg/cuBlas/Mgcg/ConjugateGradient.cs:56-79.
 * returns 0 = continue, 1 = converged, -1 = "throw" (Iteration > MaxIteration).                  */
static int ref_is_converged(int iteration, int min_iteration, int max_iteration, double residual, double allowable)
{
    if (iteration < min_iteration) return 0;
    else if (iteration > max_iteration) return -1;
    return residual < allowable;
}

/*
 * ConjugateGradientCpu.Solve — Mgcg/cuBlas/Mgcg/ConjugateGradientCpu.cs:45-98 (norm = ABS_L2).
 *   x is the initial guess on entry, the solution on exit (x0 semantics of Mgcg.cu:225).
 *   *iteration = 0-based index of the converging iteration (the C# `Iteration`); the native exports
 *   report index+1 (Mgcg.cu:234 `iteration++` after the converged pass; ConjugateGradientSingleGpu.cs:168
 *   subtracts it again).  *residual = sqrt(r.r) (or the mode's measure).
 * norm = ABS_LINF follows Mgcg/HandmadeCL/MgcgCL/ConjugateGradientCpu.cs:57-95 (rr recomputed per pass,
 *   Residual = max|r_i|); norm = REL_L2 follows Mgcg/ViennaCL/Mgcg/ComputerCpu.cpp:53-97
 *   (rrNew/rr0 < tol^2, reported residual = sqrt(rrNew/rr0)).
 * min_strict != 0 uses the SimpleConjugateGradient.cu:107 / ComputerCpu.cpp:82 test `min < iteration`
 *   instead of the C# `!(iteration < min)`; max_iteration is honoured like the C# path (returns
 *   REF_MAX_ITERATION_EXCEEDED where the C# throws).
 * Work arrays r, p, Ap are caller-provided (length n) like the reference (ConjugateGradientCpu.cs:37-39).
 */
int ref_cg_solve(int n, const double *elements, const int *row_offsets, const int *column_indeces,
                 double *x, const double *b, double *r, double *p, double *Ap,
                 double allowable_residual, int min_iteration, int max_iteration,
                 int norm, int min_strict, int *iteration, double *residual)
{
    ref_spmv(n, elements, row_offsets, column_indeces, x, Ap);          /* :54 */
    ref_set_added(n, r, b, Ap, -1);                                     /* :55 */
    memcpy(p, r, sizeof(double) * (size_t)n);                           /* :56 */
    double rr = ref_dot(n, r, r);                                       /* :57 */
    const double rr0 = rr;
    int it;
    for (it = 0;; it++) {                                               /* :60 */
        if (norm == REF_NORM_ABS_LINF) rr = ref_dot(n, r, r);           /* HandmadeCL CPU :68 */
        ref_spmv(n, elements, row_offsets, column_indeces, p, Ap);      /* :70 */
        double alpha = rr / ref_dot(n, p, Ap);                          /* :71 */
        ref_set_added(n, x, x, p, alpha);                               /* :72 */
        ref_set_added(n, r, r, Ap, -alpha);                             /* :73 */
        double rrNew = ref_dot(n, r, r);                                /* :74 */
        double res, allow = allowable_residual;
        if (norm == REF_NORM_ABS_L2) res = sqrt(rrNew);                 /* :77 */
        else if (norm == REF_NORM_ABS_LINF) res = ref_max_absolute(n, r);
        else { res = sqrt(rrNew / rr0); }                               /* ComputerCpu.cpp:79,82 (same test, sqrt both sides) */
        *residual = res;
        /* the measure-vs-tolerance test of the mode; REL_L2 is tested squared like ComputerCpu.cpp:82 */
        int below = (norm == REF_NORM_REL_L2) ? (rrNew / rr0 < allow * allow) : (res < allow);
        int c;
        if (min_strict) {                                               /* SimpleConjugateGradient.cu:107 */
            if (it > max_iteration) c = -1;
            else c = (min_iteration < it) && below;
        } else {                                                        /* :80 -> ConjugateGradient.cs:64-77 */
            c = ref_is_converged(it, min_iteration, max_iteration, below ? 0.0 : 1.0, 0.5);
        }
        if (c < 0) { *iteration = it; return REF_MAX_ITERATION_EXCEEDED; }
        if (c) break;                                                   /* :83 */
        double beta = rrNew / rr;                                       /* :93 */
        ref_set_added(n, p, r, p, beta);                                /* :94 */
        rr = rrNew;                                                     /* :95 */
    }
    *iteration = it;
    return REF_OK;
}

/* ---- multi-GPU partition rules (bit-exact integer maps) ---------------------------------------- */

/* ConjugateGradientParallelGpu ctor — Mgcg/cuBlas/Mgcg/ConjugateGradientParallelGpu.cs:271-277:
 * offsets[0]=0; offsets[i]=offsets[i-1]+floor(Count/deviceCount) (1<=i<D); offsets[D]=Count.         */
void ref_partition_offsets(int count, int device_count, int *offsets /* [device_count+1] */)
{
    offsets[0] = 0;
    for (int i = 1; i < device_count; i++)
        offsets[i] = offsets[i - 1] + (int)floor((double)count / device_count);
    offsets[device_count] = count;
}

/* Initialize — Mgcg/cuBlas/MgcgGpu/Mgcg.cu:57-85 + caller ConjugateGradientParallelGpu.cs:364-377.
 * Produces the device-local row offsets (rebased by elementOffsetForDevice, Mgcg.cu:73), the element
 * slice bounds and minJ/maxJ (Mgcg.cu:83-84).  local_row_offsets has countForDevice+1 entries.       */
void ref_initialize_block(const int *row_offsets, const int *column_indeces,
                          int offset_for_device, int count_for_device,
                          int *local_row_offsets, int *element_offset, int *element_count,
                          int *min_j, int *max_j)
{
    int eo = row_offsets[offset_for_device];                                        /* .cs:367 */
    int ec = row_offsets[offset_for_device + count_for_device] - eo;                /* .cs:366 */
    for (int i = 0; i <= count_for_device; i++)
        local_row_offsets[i] = row_offsets[offset_for_device + i] - eo;             /* Mgcg.cu:70,73 */
    int mn = column_indeces[eo], mx = column_indeces[eo];
    for (int k = eo; k < eo + ec; k++) {
        if (column_indeces[k] < mn) mn = column_indeces[k];
        if (column_indeces[k] > mx) mx = column_indeces[k];
    }
    *element_offset = eo; *element_count = ec; *min_j = mn; *max_j = mx;
}

/* Halo widths — ConjugateGradientParallelGpu.cs:397-398. */
void ref_halo_counts(int device_id, int device_count, int offset_for_device, int count_for_device,
                     int min_j, int max_j, int *last_count, int *next_count)
{
    *last_count = (device_id > 0) ? offset_for_device - min_j : 0;
    *next_count = (device_id < device_count - 1) ? max_j - count_for_device - offset_for_device + 1 : 0;
}

/*
 * ConjugateGradientParallelGpu.Solve — ConjugateGradientParallelGpu.cs:424-565 with the native phases
 * Solve0..3 (Mgcg.cu:116-198) and SyncP (cs:384-419 -> Mgcg.cu:88-113) executed sequentially per device on
 * the CPU.  Every device keeps a full-length, global-indexed p (cs:321) of which only its own block and
 * the halo ranges are ever valid; partial dots are summed in device order (cs:463,499,525 `resultsDot.Sum()`).
 * Device-level dots use the sequential ref_dot (the reference's cuBLAS order is closed-source/unpinned).
 */
int ref_cg_solve_parallel(int n, int device_count,
                          const double *elements, const int *row_offsets, const int *column_indeces,
                          double *x, const double *b,
                          double allowable_residual, int min_iteration, int max_iteration,
                          int *iteration, double *residual)
{
    int D = device_count, status = REF_OK;
    int *off = (int *)malloc(sizeof(int) * (size_t)(D + 1));
    int *minJ = (int *)malloc(sizeof(int) * (size_t)D), *maxJ = (int *)malloc(sizeof(int) * (size_t)D);
    int *eoff = (int *)malloc(sizeof(int) * (size_t)D), *ecnt = (int *)malloc(sizeof(int) * (size_t)D);
    int **lro = (int **)malloc(sizeof(int *) * (size_t)D);
    double **P = (double **)malloc(sizeof(double *) * (size_t)D);
    double *r = (double *)calloc((size_t)n, sizeof(double)), *Ap = (double *)calloc((size_t)n, sizeof(double));
    double *host = (double *)calloc((size_t)n, sizeof(double));   /* bufferHost, cs:282 */
    double *dots = (double *)malloc(sizeof(double) * (size_t)D);
    ref_partition_offsets(n, D, off);
    for (int d = 0; d < D; d++) {
        int cnt = off[d + 1] - off[d];
        lro[d] = (int *)malloc(sizeof(int) * (size_t)(cnt + 1));
        ref_initialize_block(row_offsets, column_indeces, off[d], cnt, lro[d], &eoff[d], &ecnt[d], &minJ[d], &maxJ[d]);
        P[d] = (double *)calloc((size_t)n, sizeof(double));
        memcpy(P[d] + off[d], x + off[d], sizeof(double) * (size_t)cnt);            /* Mgcg.cu:80 */
    }
#define SYNC_P()                                                                                          \
    do {                                                                                                  \
        for (int d = 0; d < D; d++) {                                                                     \
            int cnt = off[d + 1] - off[d], lc, nc;                                                        \
            ref_halo_counts(d, D, off[d], cnt, minJ[d], maxJ[d], &lc, &nc);                               \
            memcpy(host + off[d], P[d] + off[d], sizeof(double) * (size_t)lc);             /* :97 */     \
            memcpy(host + off[d] + cnt - nc, P[d] + off[d] + cnt - nc, sizeof(double) * (size_t)nc);      \
        }                                                                                                 \
        for (int d = 0; d < D; d++) {                                                                     \
            int cnt = off[d + 1] - off[d], lc, nc;                                                        \
            ref_halo_counts(d, D, off[d], cnt, minJ[d], maxJ[d], &lc, &nc);                               \
            memcpy(P[d] + off[d] - lc, host + off[d] - lc, sizeof(double) * (size_t)lc);   /* :111 */    \
            memcpy(P[d] + off[d] + cnt, host + off[d] + cnt, sizeof(double) * (size_t)nc); /* :112 */    \
        }                                                                                                 \
    } while (0)
#define SUM_DOTS(out) do { double s_ = 0; for (int d = 0; d < D; d++) s_ += dots[d]; (out) = s_; } while (0)

    SYNC_P();                                                                       /* cs:427 */
    for (int d = 0; d < D; d++) {                                                   /* Solve0, Mgcg.cu:138-141 */
        int cnt = off[d + 1] - off[d];
        ref_spmv(cnt, elements + eoff[d], lro[d], column_indeces + eoff[d], P[d], Ap + off[d]);
        ref_set_added(cnt, r + off[d], b + off[d], Ap + off[d], -1);
        memcpy(P[d] + off[d], r + off[d], sizeof(double) * (size_t)cnt);
        dots[d] = ref_dot(cnt, r + off[d], r + off[d]);
    }
    double rr; SUM_DOTS(rr);                                                        /* cs:463 */
    int it;
    for (it = 0;; it++) {                                                           /* cs:466 */
        SYNC_P();                                                                   /* cs:469 */
        for (int d = 0; d < D; d++) {                                               /* Solve1, Mgcg.cu:161-162 */
            int cnt = off[d + 1] - off[d];
            ref_spmv(cnt, elements + eoff[d], lro[d], column_indeces + eoff[d], P[d], Ap + off[d]);
            dots[d] = ref_dot(cnt, P[d] + off[d], Ap + off[d]);
        }
        double pAp; SUM_DOTS(pAp);
        double alpha = rr / pAp;                                                    /* cs:499 */
        for (int d = 0; d < D; d++) {                                               /* Solve2, Mgcg.cu:181-183 */
            int cnt = off[d + 1] - off[d];
            ref_set_added(cnt, x + off[d], x + off[d], P[d] + off[d], alpha);
            ref_set_added(cnt, r + off[d], r + off[d], Ap + off[d], -alpha);
            dots[d] = ref_dot(cnt, r + off[d], r + off[d]);
        }
        double rrNew; SUM_DOTS(rrNew);                                              /* cs:525 */
        *residual = sqrt(rrNew);                                                    /* cs:528 */
        int c = ref_is_converged(it, min_iteration, max_iteration, *residual, allowable_residual);
        if (c < 0) { status = REF_MAX_ITERATION_EXCEEDED; break; }
        if (c) break;
        double beta = rrNew / rr;                                                   /* cs:544 */
        for (int d = 0; d < D; d++) {                                               /* Solve3, Mgcg.cu:197: p*=beta; p+=r */
            int cnt = off[d + 1] - off[d];
            double *pd = P[d] + off[d];
            for (int i = 0; i < cnt; i++) { pd[i] = pd[i] * beta; pd[i] = pd[i] + r[off[d] + i]; }
        }
        rr = rrNew;
    }
    *iteration = it;
    for (int d = 0; d < D; d++) { free(lro[d]); free(P[d]); }
    free(off); free(minJ); free(maxJ); free(eoff); free(ecnt); free(lro); free(P); free(r); free(Ap); free(host); free(dots);
    return status;
#undef SYNC_P
#undef SUM_DOTS
}

/* ---- the reference's two input generators ------------------------------------------------------ */

/* SimpleConjugateGradient main — SimpleConjugateGradient/SimpleConjugateGradient/SimpleConjugateGradient.cu:163-197.
 * Tridiagonal [1 2 1], diagonal stored first; b[i] = i*i*0.5 with i*i evaluated in 32-bit int (wraps for i>=46341).
 * Arrays: elements/column_indeces [3n], row_offsets [n+1], b [n].  Returns nnz.                         */
int ref_gen_tridiag121(int n, double *elements, int *column_indeces, int *row_offsets, double *b)
{
    int nz = 0;
    row_offsets[0] = 0;
    for (int i = 0; i < n; i++) {
        elements[nz] = 2; column_indeces[nz] = i; nz++;
        if (i > 0) { elements[nz] = 1; column_indeces[nz] = i - 1; nz++; }
        if (i < n - 1) { elements[nz] = 1; column_indeces[nz] = i + 1; nz++; }
        row_offsets[i + 1] = nz;
    }
    for (int i = 0; i < n; i++) {
        int32_t ii = (int32_t)((uint32_t)i * (uint32_t)i);   /* int*int wrap, :196 */
        b[i] = ii * 0.5;
    }
    return nz;
}

/* MgcgMain.Main — Mgcg/cuBlas/Mgcg/MgcgMain.cs:51-104.  a_ij=|sin(i+j)| for j in [max(0,i-M/2+1), min(N,i+M/2)),
 * diagonal (stored FIRST in the row) = sum of the off-diagonals; b_i = 10 cos(i); x0_i = i/100.
 * Arrays sized n*max_nonzero like SparseMatrix.cs:33-34.  Returns nnz.                                  */
int ref_gen_banded_sin(int n, int max_nonzero, double *elements, int *column_indeces, int *row_offsets,
                       double *b, double *x0)
{
    row_offsets[0] = 0;
    for (int i = 0; i < n; i++) {
        int ro = row_offsets[i];
        elements[ro] = 0; column_indeces[ro] = i;
        int cnt = 1;
        int j0 = i - max_nonzero / 2 + 1; if (j0 < 0) j0 = 0;
        int j1 = i + max_nonzero / 2; if (j1 > n) j1 = n;
        for (int j = j0; j < j1; j++) {
            if (i != j) {
                double a = fabs(sin((double)(i + j)));
                elements[ro + cnt] = a; column_indeces[ro + cnt] = j; cnt++;
                elements[ro] += a;
            }
        }
        row_offsets[i + 1] = ro + cnt;
    }
    for (int i = 0; i < n; i++) { b[i] = cos((double)i) * 10; x0[i] = (double)i / 100; }
    return row_offsets[n];
}

/* ---- the OpenCL variant's ELL container and solver (Mgcg/HandmadeCL/MgcgCL) --------------------------- */

/* SparseMatrix.Multiply — Mgcg/HandmadeCL/MgcgCL/SparseMatrix.cs:202-224: answer[i] = 0, then the row's slots
 * k = 0 .. NonzeroCounts[i]-1 in order (slot 0 is the diagonal, :74-92), element index i*MaxNonzeroCountPerRow + k. */
void ref_ell_multiply(int rows, int max_nonzero, const double *elements, const int *column_indeces,
                      const int *nonzero_counts, const double *vector, double *answer)
{
    for (int i = 0; i < rows; i++) {
        answer[i] = 0;
        for (int k = 0; k < nonzero_counts[i]; k++) {
            size_t index = (size_t)i * (size_t)max_nonzero + (size_t)k;      /* GetGlobalIndex, :191-194 */
            int j = column_indeces[index];
            answer[i] += elements[index] * vector[j];
        }
    }
}

/* Index bookkeeping only: the used slots of every row, in slot order, as CSR.  Returns the entry count. */
int ref_ell_to_csr(int rows, int max_nonzero, const double *elements, const int *column_indeces, const int *nonzero_counts,
                   double *csr_elements, int *row_offsets, int *csr_column_indeces)
{
    int o = 0;
    for (int i = 0; i < rows; i++) {
        row_offsets[i] = o;
        for (int k = 0; k < nonzero_counts[i]; k++, o++) {
            size_t index = (size_t)i * (size_t)max_nonzero + (size_t)k;
            csr_elements[o] = elements[index];
            csr_column_indeces[o] = column_indeces[index];
        }
    }
    row_offsets[rows] = o;
    return o;
}

/* ConjugateGradientSingleGpu.Solve of the OpenCL variant — Mgcg/HandmadeCL/MgcgCL/ConjugateGradientSingleGpu.cs:226-296,
 * with its kernels as the sequential loops they parallelise (Mgcg.cl:15-33 result = left + C*right; :35-108 dot by
 * multiply + tree sum, restated left to right; :110-160 max |.|).  rr is recomputed at the top of every pass (:261),
 * Residual = max|r_i| (:268), IsConverged as ConjugateGradient.cs of that variant (Iteration < Min -> no;
 * Iteration > Max -> throw; Residual < Allowable).  x holds the initial guess (:217, :246). */
int ref_cg_solve_ell(int n, int max_nonzero, const double *elements, const int *column_indeces, const int *nonzero_counts,
                     double *x, const double *b, double *r, double *p, double *Ap,
                     double allowable_residual, int min_iteration, int max_iteration, int *iteration, double *residual)
{
    ref_ell_multiply(n, max_nonzero, elements, column_indeces, nonzero_counts, x, Ap);      /* :246 */
    ref_set_added(n, r, b, Ap, -1);                                                         /* :247 */
    memcpy(p, r, sizeof(double) * (size_t)n);                                               /* :248 */
    int it;
    for (it = 0;; it++) {                                                                   /* :251 */
        double rr = ref_dot(n, r, r);                                                       /* :261 */
        ref_ell_multiply(n, max_nonzero, elements, column_indeces, nonzero_counts, p, Ap);  /* :262 */
        double alpha = rr / ref_dot(n, p, Ap);                                              /* :263 */
        ref_set_added(n, x, x, p, alpha);                                                   /* :264 */
        ref_set_added(n, r, r, Ap, -alpha);                                                 /* :265 */
        *residual = ref_max_absolute(n, r);                                                 /* :268 */
        int c = ref_is_converged(it, min_iteration, max_iteration, *residual, allowable_residual);
        if (c < 0) { *iteration = it; return REF_MAX_ITERATION_EXCEEDED; }
        if (c) break;                                                                       /* :271-275 */
        double rrNew = ref_dot(n, r, r);                                                    /* :285 */
        double beta = rrNew / rr;                                                           /* :286 */
        ref_set_added(n, p, r, p, beta);                                                    /* :287 */
    }
    *iteration = it;
    return REF_OK;
}
