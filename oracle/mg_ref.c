/*
 * oracle/mg_ref.c — TEST INFRASTRUCTURE ONLY: CPU twin of the multigrid-preconditioned CG.
 *
 * PARITY UNPINNED: the reference contains NO multigrid and no preconditioner (SURVEY.md F1 — "Mgcg" appears only
 * in two doc-comments, Mgcg/cuBlas/Mgcg/MgcgMain.cs:8; the planned `p_0 = (LDLr)_0` of
 * Mgcg/HandmadeCL/MgcgCL/ConjugateGradientCpu.cs:51 was never written).  There is no reference output to pin this
 * file against.  It is a sequential restatement of the SAME algorithm the CUDA path runs (aggregation, prolongator
 * smoothing, Galerkin products with ordered summation, Gauss-Jordan coarse inverse, V-cycle, PCG in the operation
 * order of ConjugateGradientCpu.cs:45-98 with z = M r inserted), cross-checked in tests against the unpreconditioned
 * oracle (cg_ref.c) and scipy direct solves.  Compile with -ffp-contract=off.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

void ref_spmv(int rows, const double *elements, const int *row_offsets, const int *column_indeces, const double *vector, double *answer);
double ref_dot(int n, const double *left, const double *right);
double ref_max_absolute(int n, const double *v);

typedef struct { int rows, cols, nnz; double *val; int *rowptr; int *col; } csr_t;
typedef struct {
    csr_t A, P, R;
    int *agg;
    double *dinv, *x, *x2, *b, *r;
    int dims[3];
    int ownA;
    int rb;      /* smoothed by red-black Gauss-Seidel (mg_ref_set_smoother) */
    int cheb;    /* Chebyshev polynomial smoother of this degree in D^-1 A (0: off; mg_ref_set_chebyshev) */
    double lmax, lmin;   /* its eigenvalue window: lmax = max_i sum_j |a_ij| / |a_ii| (Gershgorin), lmin = lmax / ratio */
    double *d;   /* its direction vector */
} level_t;
typedef struct { int nl; level_t *L; double *coarse_inv; int coarse_n; } mg_ref_t;

static void csr_free(csr_t *m) { free(m->val); free(m->rowptr); free(m->col); memset(m, 0, sizeof(*m)); }

typedef struct { int col; int ord; double v; } ent_t;
static int ent_cmp(const void *a, const void *b)
{
    const ent_t *x = (const ent_t *)a, *y = (const ent_t *)b;
    if (x->col != y->col) return x->col < y->col ? -1 : 1;
    return x->ord < y->ord ? -1 : (x->ord > y->ord);
}

/* growable CSR builder: append one row given its entries in EXPANSION order; duplicates (same col) are summed
 * sequentially in that order after a stable sort by col == the device's stable radix sort + ordered segment sum */
typedef struct { csr_t m; size_t cap; } builder_t;
static void b_init(builder_t *b, int rows, int cols)
{
    memset(b, 0, sizeof(*b));
    b->m.rows = rows; b->m.cols = cols;
    b->m.rowptr = (int *)calloc((size_t)rows + 1, sizeof(int));
    b->cap = 1024;
    b->m.val = (double *)malloc(b->cap * sizeof(double));
    b->m.col = (int *)malloc(b->cap * sizeof(int));
}
static void b_row(builder_t *b, int row, ent_t *e, int cnt)
{
    for (int i = 0; i < cnt; i++) e[i].ord = i;
    qsort(e, (size_t)cnt, sizeof(ent_t), ent_cmp);
    int i = 0;
    while (i < cnt) {
        double s = e[i].v;
        int j = i + 1;
        while (j < cnt && e[j].col == e[i].col) { s = s + e[j].v; j++; }
        if ((size_t)b->m.nnz + 1 > b->cap) {
            b->cap *= 2;
            b->m.val = (double *)realloc(b->m.val, b->cap * sizeof(double));
            b->m.col = (int *)realloc(b->m.col, b->cap * sizeof(int));
        }
        b->m.val[b->m.nnz] = s; b->m.col[b->m.nnz] = e[i].col; b->m.nnz++;
        i = j;
    }
    b->m.rowptr[row + 1] = b->m.nnz;
}

/* C = A*B, expansion order: A's entries of the row in storage order, each times B's row in storage order */
static csr_t spgemm(const csr_t *A, const csr_t *B)
{
    builder_t bd; b_init(&bd, A->rows, B->cols);
    size_t cap = 256; ent_t *e = (ent_t *)malloc(cap * sizeof(ent_t));
    for (int i = 0; i < A->rows; i++) {
        int cnt = 0;
        for (int k = A->rowptr[i]; k < A->rowptr[i + 1]; k++) {
            int j = A->col[k]; double a = A->val[k];
            for (int q = B->rowptr[j]; q < B->rowptr[j + 1]; q++) {
                if ((size_t)cnt + 1 > cap) { cap *= 2; e = (ent_t *)realloc(e, cap * sizeof(ent_t)); }
                e[cnt].col = B->col[q]; e[cnt].v = a * B->val[q]; cnt++;
            }
        }
        b_row(&bd, i, e, cnt);
    }
    free(e);
    return bd.m;
}

/* T = A^T; within a row of T entries are in ascending original-row order (== sort by (col,row)) */
static csr_t transpose(const csr_t *A)
{
    csr_t T; T.rows = A->cols; T.cols = A->rows; T.nnz = A->nnz;
    T.rowptr = (int *)calloc((size_t)T.rows + 1, sizeof(int));
    T.val = (double *)malloc(sizeof(double) * (size_t)(A->nnz > 0 ? A->nnz : 1));
    T.col = (int *)malloc(sizeof(int) * (size_t)(A->nnz > 0 ? A->nnz : 1));
    for (int k = 0; k < A->nnz; k++) T.rowptr[A->col[k] + 1]++;
    for (int i = 0; i < T.rows; i++) T.rowptr[i + 1] += T.rowptr[i];
    int *fill = (int *)malloc(sizeof(int) * (size_t)(T.rows > 0 ? T.rows : 1));
    memcpy(fill, T.rowptr, sizeof(int) * (size_t)T.rows);
    for (int i = 0; i < A->rows; i++)
        for (int k = A->rowptr[i]; k < A->rowptr[i + 1]; k++) { int p = fill[A->col[k]]++; T.col[p] = i; T.val[p] = A->val[k]; }
    free(fill);
    return T;
}

/* in-place Gauss-Jordan inverse, same update formulas and order as k_gj_pivot / k_gj_update */
static double *dense_inverse(const csr_t *A)
{
    int n = A->rows;
    double *M = (double *)calloc((size_t)n * n, sizeof(double));
    double *colk = (double *)malloc(sizeof(double) * (size_t)n), *rowk = (double *)malloc(sizeof(double) * (size_t)n);
    for (int i = 0; i < n; i++) for (int k = A->rowptr[i]; k < A->rowptr[i + 1]; k++) M[(size_t)i * n + A->col[k]] = A->val[k];
    for (int k = 0; k < n; k++) {
        double p = M[(size_t)k * n + k];
        for (int j = 0; j < n; j++) { colk[j] = M[(size_t)j * n + k]; rowk[j] = (j == k) ? 1.0 / p : M[(size_t)k * n + j] / p; }
        for (int i = 0; i < n; i++)
            for (int j = 0; j < n; j++) {
                double v;
                if (i == k) v = rowk[j];
                else if (j == k) v = -(colk[i] * rowk[k]);
                else v = M[(size_t)i * n + j] + (-(colk[i] * rowk[j]));
                M[(size_t)i * n + j] = v;
            }
    }
    free(colk); free(rowk);
    return M;
}


/* ---- strength-of-connection filter + MIS-2 aggregation for matrices without a grid hint (device: csrc/mg_aggregate.cu).
 * Every step is a synchronous (Jacobi-style) sweep over the rows so that the sequential twin and the one-thread-per-row
 * device kernels produce the same integers.
 *   rowmax_i   = max_{col_k != i} |a_ik|
 *   strong(i,k): col_k != i, a_ik != 0, |a_ik| >= theta * min(rowmax_i, rowmax_col_k)      (symmetric for symmetric A)
 *   filtered matrix A_F: row i = [ (i, d_F) , the strong entries in storage order ], d_F = (sum of the diagonal entries)
 *                + (the weak entries, added in storage order) when that stays positive, else the diagonal alone.
 *   MIS-2 of the strong graph by rounds: T = (state, hash(i), i) packed in 64 bits; two max-propagation hops; an
 *                undecided node that is the maximum of its 2-hop neighbourhood joins the set, one that sees a member
 *                leaves it.   Roots are numbered in index order; two passes attach every other node to the aggregate of
 *                the strong neighbour with the largest |a_ij| (first in storage order on ties); leftovers are singletons. */
static uint32_t mix32(uint32_t x)
{
    x *= 0x9E3779B1u; x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
    return x;
}

static csr_t strength_filter(const csr_t *A, double theta, double **dinvF_out)
{
    int n = A->rows;
    double *rowmax = (double *)calloc((size_t)(n > 0 ? n : 1), sizeof(double));
    for (int i = 0; i < n; i++) {
        double m = 0.0;
        for (int k = A->rowptr[i]; k < A->rowptr[i + 1]; k++)
            if (A->col[k] != i) { double a = fabs(A->val[k]); if (a > m) m = a; }
        rowmax[i] = m;
    }
    csr_t F; F.rows = n; F.cols = n;
    F.rowptr = (int *)calloc((size_t)n + 1, sizeof(int));
    for (int i = 0; i < n; i++) {
        int cnt = 1;
        for (int k = A->rowptr[i]; k < A->rowptr[i + 1]; k++) {
            int j = A->col[k];
            if (j == i) continue;
            double a = A->val[k], mm = rowmax[i] < rowmax[j] ? rowmax[i] : rowmax[j];
            if (a != 0.0 && fabs(a) >= theta * mm) cnt++;
        }
        F.rowptr[i + 1] = F.rowptr[i] + cnt;
    }
    F.nnz = F.rowptr[n];
    F.val = (double *)malloc(sizeof(double) * (size_t)(F.nnz > 0 ? F.nnz : 1));
    F.col = (int *)malloc(sizeof(int) * (size_t)(F.nnz > 0 ? F.nnz : 1));
    double *dinvF = (double *)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; i++) {
        int o = F.rowptr[i] + 1;
        double d = 0.0;
        for (int k = A->rowptr[i]; k < A->rowptr[i + 1]; k++) if (A->col[k] == i) d = d + A->val[k];
        double l = d;
        for (int k = A->rowptr[i]; k < A->rowptr[i + 1]; k++) {
            int j = A->col[k];
            if (j == i) continue;
            double a = A->val[k], mm = rowmax[i] < rowmax[j] ? rowmax[i] : rowmax[j];
            if (a != 0.0 && fabs(a) >= theta * mm) { F.val[o] = a; F.col[o] = j; o++; }
            else l = l + a;
        }
        double dF = (l > 0.0) ? l : d;
        if (dF == 0.0) dF = 1.0;
        F.val[F.rowptr[i]] = dF; F.col[F.rowptr[i]] = i;
        dinvF[i] = 1.0 / dF;
    }
    free(rowmax);
    *dinvF_out = dinvF;
    return F;
}

/* F: filtered matrix (diagonal first).  Returns the number of aggregates; agg[i] filled. */
static int mis2_aggregate(const csr_t *F, int *agg)
{
    int n = F->rows;
    size_t nn = (size_t)(n > 0 ? n : 1);
    uint64_t *T0 = (uint64_t *)malloc(8 * nn), *T1 = (uint64_t *)malloc(8 * nn), *T2 = (uint64_t *)malloc(8 * nn);
    unsigned char *state = (unsigned char *)malloc(nn);
    for (int i = 0; i < n; i++) state[i] = 1;
    for (int round = 0; round < 256; round++) {
        int undecided = 0;
        for (int i = 0; i < n; i++) {
            T0[i] = ((uint64_t)state[i] << 62) | ((uint64_t)(mix32((uint32_t)i) >> 2) << 32) | (uint32_t)i;
            undecided += state[i] == 1;
        }
        if (!undecided) break;
        for (int i = 0; i < n; i++) {
            uint64_t m = T0[i];
            for (int k = F->rowptr[i] + 1; k < F->rowptr[i + 1]; k++) if (T0[F->col[k]] > m) m = T0[F->col[k]];
            T1[i] = m;
        }
        for (int i = 0; i < n; i++) {
            uint64_t m = T1[i];
            for (int k = F->rowptr[i] + 1; k < F->rowptr[i + 1]; k++) if (T1[F->col[k]] > m) m = T1[F->col[k]];
            T2[i] = m;
        }
        for (int i = 0; i < n; i++)
            if (state[i] == 1) {
                if (T2[i] == T0[i]) state[i] = 2;
                else if ((T2[i] >> 62) == 2) state[i] = 0;
            }
    }
    int nroots = 0;
    for (int i = 0; i < n; i++) agg[i] = (state[i] == 2) ? nroots++ : -1;
    int *src = (int *)malloc(sizeof(int) * nn);
    for (int pass = 0; pass < 2; pass++) {
        memcpy(src, agg, sizeof(int) * nn);
        for (int i = 0; i < n; i++) {
            if (src[i] >= 0) continue;
            double best = -1.0; int pick = -1;
            for (int k = F->rowptr[i] + 1; k < F->rowptr[i + 1]; k++) {
                int j = F->col[k];
                if (src[j] < 0) continue;
                double a = fabs(F->val[k]);
                if (a > best) { best = a; pick = src[j]; }
            }
            agg[i] = pick;
        }
    }
    int nc = nroots;
    for (int i = 0; i < n; i++) if (agg[i] < 0) agg[i] = nc++;
    free(src); free(state); free(T0); free(T1); free(T2);
    return nc;
}

static void level_vectors(level_t *L)
{
    size_t n = (size_t)L->A.rows;
    L->x = (double *)calloc(n, sizeof(double)); L->x2 = (double *)calloc(n, sizeof(double));
    L->b = (double *)calloc(n, sizeof(double)); L->r = (double *)calloc(n, sizeof(double));
    L->dinv = (double *)malloc(n * sizeof(double));
    for (int i = 0; i < L->A.rows; i++) {
        double d = 1.0;
        for (int k = L->A.rowptr[i]; k < L->A.rowptr[i + 1]; k++) if (L->A.col[k] == i) { d = L->A.val[k]; break; }
        L->dinv[i] = 1.0 / d;
    }
}

mg_ref_t *mg_ref_build(int n, int nnz, const double *val, const int *rowptr, const int *col,
                       int nx, int ny, int nz, int max_levels, int coarse_size, int smoothed, double prolong_omega,
                       int smoothed_levels, int agg_size, int agg_size_coarse, int aggregation, double theta)
{
    mg_ref_t *H = (mg_ref_t *)calloc(1, sizeof(mg_ref_t));
    if (max_levels <= 0) max_levels = 12;
    if (coarse_size <= 0) coarse_size = 512;
    H->L = (level_t *)calloc((size_t)max_levels + 1, sizeof(level_t));
    level_t *L0 = &H->L[0];
    L0->A.rows = n; L0->A.cols = n; L0->A.nnz = nnz;
    L0->A.val = (double *)val; L0->A.rowptr = (int *)rowptr; L0->A.col = (int *)col; L0->ownA = 0;
    int hint = (nx > 0 && ny > 0 && nz > 0 && (long long)nx * ny * nz == n);
    L0->dims[0] = hint ? nx : 0; L0->dims[1] = hint ? ny : 0; L0->dims[2] = hint ? nz : 0;
    H->nl = 1;
    for (int l = 0;; l++) {
        level_t *L = &H->L[l];
        int m = L->A.rows;
        level_vectors(L);
        if (m <= coarse_size || l + 1 >= max_levels) break;
        int nc, cd[3] = {0, 0, 0};
        L->agg = (int *)malloc(sizeof(int) * (size_t)m);
        if (L->dims[0] > 0 && aggregation != 2) {
            int a = (l == 0 || agg_size_coarse <= 0) ? agg_size : agg_size_coarse;
            if (a < 2) a = 2;
            cd[0] = (L->dims[0] + a - 1) / a; cd[1] = (L->dims[1] + a - 1) / a; cd[2] = (L->dims[2] + a - 1) / a;
            nc = cd[0] * cd[1] * cd[2];
            for (int i = 0; i < m; i++) {
                int x = i % L->dims[0], y = (i / L->dims[0]) % L->dims[1], z = i / (L->dims[0] * L->dims[1]);
                L->agg[i] = (x / a) + cd[0] * ((y / a) + cd[1] * (z / a));
            }
        } else if (aggregation == 1) {
            nc = (m + 7) / 8;
            for (int i = 0; i < m; i++) L->agg[i] = i / 8;
        } else {
            nc = -1;   /* MIS-2 below */
        }
        /* aggregation: 0 auto (geometric with a hint, MIS-2 without), 1 blocks of 8 rows without a hint, 2 MIS-2 always */
        csr_t F; memset(&F, 0, sizeof(F));
        double *dinvF = NULL;
        const int filtered = (aggregation == 2) || (L->dims[0] <= 0 && aggregation != 1);
        if (filtered) {
            F = strength_filter(&L->A, theta, &dinvF);
            nc = mis2_aggregate(&F, L->agg);
            cd[0] = cd[1] = cd[2] = 0;
        }
        if (nc >= m) { csr_free(&F); free(dinvF); break; }
        /* prolongator: smoothed with the level operator, or with its strength-filtered version under MIS-2 */
        const csr_t *Asm = filtered ? &F : &L->A;
        const double *dsm = filtered ? dinvF : L->dinv;
        builder_t bp; b_init(&bp, m, nc);
        size_t cap = 256; ent_t *e = (ent_t *)malloc(cap * sizeof(ent_t));
        for (int i = 0; i < m; i++) {
            int cnt = 0;
            e[cnt].col = L->agg[i]; e[cnt].v = 1.0; cnt++;
            if (smoothed && (smoothed_levels <= 0 || l < smoothed_levels)) {
                double f = -(prolong_omega * dsm[i]);
                for (int k = Asm->rowptr[i]; k < Asm->rowptr[i + 1]; k++) {
                    if ((size_t)cnt + 1 > cap) { cap *= 2; e = (ent_t *)realloc(e, cap * sizeof(ent_t)); }
                    e[cnt].col = L->agg[Asm->col[k]]; e[cnt].v = f * Asm->val[k]; cnt++;
                }
            }
            b_row(&bp, i, e, cnt);
        }
        free(e);
        csr_free(&F); free(dinvF);
        L->P = bp.m;
        L->R = transpose(&L->P);
        csr_t AP = spgemm(&L->A, &L->P);
        csr_t Ac = spgemm(&L->R, &AP);
        csr_free(&AP);
        level_t *N = &H->L[l + 1];
        N->A = Ac; N->ownA = 1;
        N->dims[0] = cd[0]; N->dims[1] = cd[1]; N->dims[2] = cd[2];
        H->nl = l + 2;
    }
    level_t *Lc = &H->L[H->nl - 1];
    H->coarse_n = Lc->A.rows;
    H->coarse_inv = dense_inverse(&Lc->A);
    return H;
}

void mg_ref_free(mg_ref_t *H)
{
    if (!H) return;
    for (int l = 0; l < H->nl; l++) {
        level_t *L = &H->L[l];
        if (L->ownA) csr_free(&L->A);
        csr_free(&L->P); csr_free(&L->R);
        free(L->agg); free(L->dinv); free(L->x); free(L->x2); free(L->b); free(L->r); free(L->d);
    }
    free(H->L); free(H->coarse_inv); free(H);
}

int mg_ref_levels(const mg_ref_t *H) { return H->nl; }
void mg_ref_level_size(const mg_ref_t *H, int l, int *rows, int *nnz) { *rows = H->L[l].A.rows; *nnz = H->L[l].A.nnz; }
void mg_ref_level_get(const mg_ref_t *H, int l, double *val, int *rowptr, int *col)
{
    const csr_t *A = &H->L[l].A;
    memcpy(val, A->val, sizeof(double) * (size_t)A->nnz);
    memcpy(col, A->col, sizeof(int) * (size_t)A->nnz);
    memcpy(rowptr, A->rowptr, sizeof(int) * ((size_t)A->rows + 1));
}
void mg_ref_level_agg(const mg_ref_t *H, int l, int *agg) { memcpy(agg, H->L[l].agg, sizeof(int) * (size_t)H->L[l].A.rows); }

/* smoother 0: weighted Jacobi on every level.  1: symmetric red-black Gauss-Seidel (red then black before the coarse correction,
 * black then red after it) on every level whose rows are the cells of a known grid and whose off-diagonal entries all join cells
 * of opposite parity (ix + iy + iz) & 1 — 5-/7-point operators; weighted Jacobi on the others (device: csrc/mg_cycle.cu,
 * k_check_two_colour in csrc/mg_aggregate.cu). */
static int cell_parity(const level_t *L, int i)
{
    int nx = L->dims[0], ny = L->dims[1];
    return ((i % nx) + ((i / nx) % ny) + (i / nx / ny)) & 1;
}
void mg_ref_set_smoother(mg_ref_t *H, int smoother)
{
    for (int l = 0; l < H->nl; l++) {
        level_t *L = &H->L[l];
        L->rb = 0;
        if (smoother != 1 || L->dims[0] <= 0 || l == H->nl - 1) continue;
        int ok = 1;
        for (int i = 0; i < L->A.rows && ok; i++)
            for (int k = L->A.rowptr[i]; k < L->A.rowptr[i + 1]; k++)
                if (L->A.col[k] != i && cell_parity(L, L->A.col[k]) == cell_parity(L, i)) { ok = 0; break; }
        L->rb = ok;
    }
}
int mg_ref_level_rb(const mg_ref_t *H, int l) { return H->L[l].rb; }

/* in place: rows of parity `color` get x_i += D_ii^-1 (b_i - (A x)_i), row sums in storage order */
static void gs_half_sweep(level_t *L, int color)
{
    const csr_t *A = &L->A;
    for (int i = 0; i < A->rows; i++) {
        if (cell_parity(L, i) != color) continue;
        double s = 0.0;
        for (int k = A->rowptr[i]; k < A->rowptr[i + 1]; k++) s = s + A->val[k] * L->x[A->col[k]];
        L->x[i] = L->x[i] + 1.0 * L->dinv[i] * (L->b[i] - s);
    }
}

/* Chebyshev smoothing (opt-in mode outside the reference-parity contract; device: csrc/mg_cycle.cu cheby_smooth).
 * degree steps of the Chebyshev iteration for D^-1 A on [lmin, lmax] (Saad, Iterative Methods, alg. 12.1):
 *   theta = (lmax+lmin)/2, delta = (lmax-lmin)/2, sigma = theta/delta, rho_0 = 1/sigma
 *   d_0 = (1/theta) D^-1 (b - A x) ; x += d_0
 *   rho_k = 1/(2 sigma - rho_{k-1}) ; d_k = rho_k rho_{k-1} d_{k-1} + (2 rho_k/delta) D^-1 (b - A x) ; x += d_k          */
void mg_ref_set_chebyshev(mg_ref_t *H, int degree, double ratio)
{
    for (int l = 0; l < H->nl; l++) {
        level_t *L = &H->L[l];
        L->cheb = (l == H->nl - 1) ? 0 : degree;
        if (!L->cheb) continue;
        double g = 0.0;
        for (int i = 0; i < L->A.rows; i++) {
            double s = 0.0;
            for (int k = L->A.rowptr[i]; k < L->A.rowptr[i + 1]; k++) s = s + fabs(L->A.val[k]);
            s = s * fabs(L->dinv[i]);
            if (s > g) g = s;
        }
        L->lmax = g; L->lmin = g / ratio;
        if (!L->d) L->d = (double *)calloc((size_t)(L->A.rows > 0 ? L->A.rows : 1), sizeof(double));
    }
}
double mg_ref_level_lmax(const mg_ref_t *H, int l) { return H->L[l].lmax; }

/* zero_guess: x = 0 on entry (the first step needs no product) */
static void cheby_smooth(level_t *L, int zero_guess)
{
    const csr_t *A = &L->A;
    const int m = A->rows;
    const double theta = 0.5 * (L->lmax + L->lmin), delta = 0.5 * (L->lmax - L->lmin), sigma = theta / delta;
    double rho = 1.0 / sigma;
    const double c0 = 1.0 / theta;
    if (zero_guess) {
        for (int i = 0; i < m; i++) { L->d[i] = c0 * L->dinv[i] * L->b[i]; L->x[i] = L->d[i]; }
    } else {
        ref_spmv(m, A->val, A->rowptr, A->col, L->x, L->x2);
        for (int i = 0; i < m; i++) { L->d[i] = c0 * L->dinv[i] * (L->b[i] - L->x2[i]); L->x2[i] = L->x[i] + L->d[i]; }
        double *t = L->x; L->x = L->x2; L->x2 = t;
    }
    for (int k = 1; k < L->cheb; k++) {
        const double rho_new = 1.0 / (2.0 * sigma - rho);
        const double c1 = rho_new * rho, c2 = 2.0 * rho_new / delta;
        ref_spmv(m, A->val, A->rowptr, A->col, L->x, L->x2);
        for (int i = 0; i < m; i++) { L->d[i] = c1 * L->d[i] + c2 * L->dinv[i] * (L->b[i] - L->x2[i]); L->x2[i] = L->x[i] + L->d[i]; }
        double *t = L->x; L->x = L->x2; L->x2 = t;
        rho = rho_new;
    }
}

static void jacobi_sweep(level_t *L, double omega)
{
    const csr_t *A = &L->A;
    ref_spmv(A->rows, A->val, A->rowptr, A->col, L->x, L->x2);
    for (int i = 0; i < A->rows; i++) L->x2[i] = L->x[i] + omega * L->dinv[i] * (L->b[i] - L->x2[i]);
    double *t = L->x; L->x = L->x2; L->x2 = t;
}

/* z = M r : V(pre,post) with weighted Jacobi, zero initial guess, dense-inverse coarsest solve */
void mg_ref_apply(mg_ref_t *H, int pre, int post, double omega, const double *r, double *z)
{
    int nl = H->nl;
    memcpy(H->L[0].b, r, sizeof(double) * (size_t)H->L[0].A.rows);
    for (int l = 0; l < nl - 1; l++) {
        level_t *L = &H->L[l];
        int m = L->A.rows;
        if (L->cheb) {
            cheby_smooth(L, 1);
            for (int s = 1; s < pre; s++) cheby_smooth(L, 0);
        } else if (L->rb) {
            for (int i = 0; i < m; i++) L->x[i] = 1.0 * L->dinv[i] * L->b[i];     /* == the red half sweep from a zero guess, on every row */
            gs_half_sweep(L, 1);
            for (int s = 1; s < pre; s++) { gs_half_sweep(L, 0); gs_half_sweep(L, 1); }
        } else {
            for (int i = 0; i < m; i++) L->x[i] = omega * L->dinv[i] * L->b[i];
            for (int s = 1; s < pre; s++) jacobi_sweep(L, omega);
        }
        ref_spmv(m, L->A.val, L->A.rowptr, L->A.col, L->x, L->r);
        for (int i = 0; i < m; i++) L->r[i] = L->b[i] - L->r[i];
        ref_spmv(L->R.rows, L->R.val, L->R.rowptr, L->R.col, L->r, H->L[l + 1].b);
    }
    {
        level_t *L = &H->L[nl - 1];
        int m = L->A.rows;
        for (int i = 0; i < m; i++) {
            double s = 0;
            for (int j = 0; j < m; j++) s += H->coarse_inv[(size_t)i * m + j] * L->b[j];
            L->x[i] = s;
        }
    }
    for (int l = nl - 2; l >= 0; l--) {
        level_t *L = &H->L[l];
        int m = L->A.rows;
        ref_spmv(m, L->P.val, L->P.rowptr, L->P.col, H->L[l + 1].x, L->x2);
        for (int i = 0; i < m; i++) L->x[i] = 1.0 * L->x2[i] + 1.0 * L->x[i];
        for (int s = 0; s < post; s++) {
            if (L->cheb) cheby_smooth(L, 0);
            else if (L->rb) { gs_half_sweep(L, 1); gs_half_sweep(L, 0); }
            else jacobi_sweep(L, omega);
        }
    }
    memcpy(z, H->L[0].x, sizeof(double) * (size_t)H->L[0].A.rows);
}

/* Preconditioned CG in the reference's operation order (ConjugateGradientCpu.cs:45-98) with z = M r:
 *   alpha = rz/pAp ; x += alpha p ; r -= alpha Ap ; stop rule on r ; z = M r ; beta = rz'/rz ; p = z + beta p.
 * norm: 0 abs-L2, 1 abs-Linf, 2 rel-L2 (see cg_ref.c).  Returns 0 ok / 3 max-iteration exceeded.          */
int mg_ref_pcg(mg_ref_t *H, int pre, int post, double omega, const double *b, double *x,
               double tol, int min_iteration, int max_iteration, int norm, int min_strict,
               int *iteration, double *residual)
{
    const csr_t *A = &H->L[0].A;
    int n = A->rows, status = 0, it;
    double *r = (double *)malloc(sizeof(double) * (size_t)n), *z = (double *)malloc(sizeof(double) * (size_t)n);
    double *p = (double *)malloc(sizeof(double) * (size_t)n), *Ap = (double *)malloc(sizeof(double) * (size_t)n);
    ref_spmv(n, A->val, A->rowptr, A->col, x, Ap);
    for (int i = 0; i < n; i++) r[i] = b[i] - Ap[i];
    double rr0 = ref_dot(n, r, r);
    mg_ref_apply(H, pre, post, omega, r, z);
    memcpy(p, z, sizeof(double) * (size_t)n);
    double rz = ref_dot(n, r, z);
    for (it = 0;; it++) {
        ref_spmv(n, A->val, A->rowptr, A->col, p, Ap);
        double pAp = ref_dot(n, p, Ap);
        double alpha = (rz == 0.0) ? 0.0 : rz / pAp;
        for (int i = 0; i < n; i++) x[i] = x[i] + alpha * p[i];
        for (int i = 0; i < n; i++) r[i] = r[i] + (-alpha) * Ap[i];
        double rr = ref_dot(n, r, r), res;
        int below;
        if (norm == 0) { res = sqrt(rr); below = res < tol; }
        else if (norm == 1) { res = ref_max_absolute(n, r); below = res < tol; }
        else if (rr0 == 0.0) { res = 0; below = 1; }
        else { res = sqrt(rr / rr0); below = (rr / rr0) < tol * tol; }
        *residual = res;
        int c;
        if (min_strict) { if (it > max_iteration) c = -1; else c = (min_iteration < it) && below; }
        else { if (it < min_iteration) c = 0; else if (it > max_iteration) c = -1; else c = below; }
        if (c < 0) { status = 3; break; }
        if (c) break;
        mg_ref_apply(H, pre, post, omega, r, z);
        double rzn = ref_dot(n, r, z);
        double beta = (rz == 0.0) ? 0.0 : rzn / rz;
        for (int i = 0; i < n; i++) p[i] = z[i] + beta * p[i];
        rz = rzn;
    }
    *iteration = it;
    free(r); free(z); free(p); free(Ap);
    return status;
}
