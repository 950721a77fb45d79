// mg.cuh — multigrid hierarchy data structures (device-resident).
#pragma once
#include "common.cuh"
#include <chrono>
#include <stdlib.h>

namespace mgcg {

// an owned device CSR
struct DevCsr {
    int rows = 0, cols = 0, nnz = 0;
    double *val = nullptr;
    int *rowptr = nullptr;
    int *col = nullptr;
    Csr view() const { Csr c; c.rows = rows; c.cols = cols; c.nnz = nnz; c.val = val; c.rowptr = rowptr; c.col = col; return c; }
    void free_all();
};

struct MgLevel {
    Csr A;                 // operator of this level (level 0: the caller's arrays)
    DevCsr Aown;           // owned storage for levels >= 1
    DevCsr P, R;           // prolongation (rows x rows_coarse) and its transpose; empty on the coarsest level
    int *agg = nullptr;    // fine -> coarse aggregate id (rows entries); nullptr on the coarsest level
    double *dinv = nullptr;
    double *xa = nullptr, *xb = nullptr, *b = nullptr, *r = nullptr;   // level vectors (level 0: b is the CG residual)
    double *xcur = nullptr;
    SpmvPlan planA, planP, planR;
    CodeStore code;        // coded columns of A on levels >= 1 (level 0: the context's store)
    int dims[3] = {0, 0, 0};
    bool rb = false;       // smoothed by red-black Gauss-Seidel: grid known and every off-diagonal entry joins cells of opposite parity
    int cheb = 0;          // smoothed by a Chebyshev polynomial of this degree in D^-1 A (0: no)
    double lmax = 0.0;     // ... on [lmax / ratio, lmax], lmax = max_i sum_j |a_ij| / |a_ii|
    double *d = nullptr;   // ... its direction vector
};
// coefficients of step k (0-based) of the Chebyshev iteration on [lmax/ratio, lmax] (Saad, alg. 12.1): d_k = c1 d_{k-1} + c2 D^-1 r
inline void cheby_coeffs(double lmax, double ratio, int k, double *c1, double *c2)
{
    const double lmin = lmax / ratio;
    const double theta = 0.5 * (lmax + lmin), delta = 0.5 * (lmax - lmin), sigma = theta / delta;
    double rho = 1.0 / sigma;
    *c1 = 0.0; *c2 = 1.0 / theta;
    for (int j = 1; j <= k; ++j) {
        const double rho_new = 1.0 / (2.0 * sigma - rho);
        *c1 = rho_new * rho; *c2 = 2.0 * rho_new / delta;
        rho = rho_new;
    }
}
// max_i sum_j |a_ij| * |dinv_i| (device scalar -> host; synchronises)                                      (mg_setup.cu)
int gershgorin_lmax(mgcg_ctx *c, const Csr &A, const double *dinv, double *out);

// arguments of the persistent coarse-tail kernel (mg_tail.cu)
constexpr int kTailMaxLevels = 8;
struct TailLevel {
    Csr A, P, R;
    const double *dinv;
    double *xa, *xb, *b, *r;
    double *xfinal;              // buffer that holds the level's x when the cycle leaves it
    int lanesA, lanesP, lanesR;
};
struct TailArgs {
    int nl;                      // levels in the tail; the last one is the dense coarsest level
    TailLevel L[kTailMaxLevels];
    const double *coarse_inv;
    int coarse_n;
    double omega;
    int pre, post;
    unsigned int *barrier;       // {arrival count, generation}
};

struct Hierarchy {
    int tail_first = -1;            // first level handled by the persistent tail kernel (-1: none)
    std::vector<MgLevel> L;
    double *coarse_inv = nullptr;   // dense inverse of the coarsest operator (n x n row-major)
    int coarse_n = 0;
    double op_complexity = 1.0;
};

// ---- mg_setup.cu building blocks (also used by tests through the introspection API)
int spgemm_esc(mgcg_ctx *c, const Csr &A, const Csr &B, DevCsr *C);
// C = A * B where row j of B is selected by (A.col[k] - col_shift)                                   (mg_spgemm.cu)
int spgemm(mgcg_ctx *c, const Csr &A, const Csr &B, DevCsr *C, int col_shift = 0);
int coo_to_csr_sum(mgcg_ctx *c, int rows, int cols, size_t count, unsigned long long *keys, double *vals, DevCsr *out);
int build_prolongator(mgcg_ctx *c, const Csr &A, int row_base, const int *agg, const double *dinv, bool smooth, double omega,
                      int ncoarse, DevCsr *P);
void set_alloc_stream(cudaStream_t st);   // stream used by the build helpers' pool allocations on this host thread
int extract_dinv(mgcg_ctx *c, const Csr &A, int row_base, double *dinv);
int csr_transpose(mgcg_ctx *c, const Csr &A, DevCsr *T);
// cnt[0..rows) (cnt[rows] == 0) -> rowptr[0..rows], total in 64 bits; MGCG_ERR_UNSUPPORTED when it does not fit int32 (mg_spgemm.cu)
int scan_counts_checked(mgcg_ctx *c, const int *cnt, int rows, int *rowptr, int *nnz, const char *what);
// ---- mg_aggregate.cu: aggregation without a grid — strength filter (A -> A_F, 1/diag(A_F)) and MIS-2 aggregates on A_F
int strength_filter(mgcg_ctx *c, const Csr &A, double theta, DevCsr *F, double **dinvF);
int mis2_aggregate(mgcg_ctx *c, const Csr &F, int *agg, int *ncoarse, int *rounds_out);
// 1 when every off-diagonal entry of A (rows = cells of an nx x ny x nz grid) joins cells of opposite parity (ix+iy+iz)&1
int check_two_colourable(mgcg_ctx *c, const Csr &A, int nx, int ny, int *result);
// ---- mg_tail.cu
void mg_tail_plan(mgcg_ctx *c);
int mg_tail_run(mgcg_ctx *c, int first, double **x_out);

// MGCG_TRACE=1: wall-clock phase marks of the hierarchy build on stderr (synchronises the stream at every mark)
struct PhaseTrace {
    bool on; cudaStream_t st; std::chrono::steady_clock::time_point t0;
    PhaseTrace(cudaStream_t s) : on(getenv("MGCG_TRACE") != nullptr), st(s) { if (on) { cudaStreamSynchronize(st); t0 = std::chrono::steady_clock::now(); } }
    void mark(int level, const char *what)
    {
        if (!on) return;
        cudaStreamSynchronize(st);
        auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[mgcg setup] level %d %-12s %8.3f ms\n", level, what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

}  // namespace mgcg
