// mg_aggregate.cu — algebraic aggregation for matrices that carry no grid: strength-of-connection filter + MIS-2 aggregates,
// built on the device.  (No counterpart in the reference; the matrices this serves are the caller-supplied CSR of
// the unchanged `Solve` signature, Mgcg/cuBlas/MgcgGpu/Mgcg.cu:201-207 — e.g. the particle pressure matrix with up to 160
// entries per row, Mgcg/cuBlas/Mgcg/MgcgMain.cs:20.)  CPU twin: oracle/mg_ref.c strength_filter / mis2_aggregate — every step
// below is a synchronous sweep with one thread per row, integer results (aggregate maps) are bit-identical to the twin, the
// filtered diagonal is summed in storage order with explicit roundings.
//
//   rowmax_i    = max_{col_k != i} |a_ik|
//   strong(i,k) : col_k != i, a_ik != 0, |a_ik| >= theta * min(rowmax_i, rowmax_col_k)          (symmetric for symmetric A)
//   A_F         : row i = [(i, d_F), strong entries in storage order]; d_F = diagonal + weak entries (storage order) when that
//                 stays positive, else the diagonal alone.  The prolongator is smoothed with A_F (Vanek's filtered matrix): its
//                 rows stay short however wide the stencil is, which bounds the Galerkin products.
//   MIS-2       : T = (state, hash(i), i) packed in 64 bits, two max-propagation hops per round; an undecided row that is the
//                 maximum of its 2-hop neighbourhood becomes a root, one that sees a root drops out.
//   aggregates  : roots numbered in index order; two passes attach each remaining row to the aggregate of its strong neighbour
//                 with the largest |a_ij| (first in storage order on ties); leftovers become singletons.
#include "mg.cuh"
#include <cub/cub.cuh>

namespace mgcg {

typedef unsigned long long u64;

static inline int agrid(long long n, int sm)
{
    long long g = (n + 255) / 256;
    if (g > (long long)sm * 32) g = (long long)sm * 32;
    return (int)(g < 1 ? 1 : g);
}

__device__ __forceinline__ unsigned mix32(unsigned x)
{
    x *= 0x9E3779B1u; x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
    return x;
}

__global__ void __launch_bounds__(256) k_rowmax(Csr A, double *__restrict__ rowmax)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.rows; i += gridDim.x * blockDim.x) {
        double m = 0.0;
        for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k)
            if (A.col[k] != i) { const double a = fabs(A.val[k]); if (a > m) m = a; }
        rowmax[i] = m;
    }
}

__device__ __forceinline__ bool is_strong(double a, double rmi, double rmj, double theta)
{
    const double mm = rmi < rmj ? rmi : rmj;
    return a != 0.0 && fabs(a) >= __dmul_rn(theta, mm);
}

__global__ void __launch_bounds__(256) k_filter_count(Csr A, const double *__restrict__ rowmax, double theta, int *__restrict__ cnt)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= A.rows; i += gridDim.x * blockDim.x) {
        if (i == A.rows) { cnt[i] = 0; continue; }
        const double rmi = rowmax[i];
        int n = 1;
        for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k) {
            const int j = A.col[k];
            if (j != i && is_strong(A.val[k], rmi, rowmax[j], theta)) ++n;
        }
        cnt[i] = n;
    }
}

__global__ void __launch_bounds__(256) k_filter_fill(Csr A, const double *__restrict__ rowmax, double theta, const int *__restrict__ f_rowptr,
                                                     double *__restrict__ f_val, int *__restrict__ f_col, double *__restrict__ dinvF)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.rows; i += gridDim.x * blockDim.x) {
        const double rmi = rowmax[i];
        const int ks = A.rowptr[i], ke = A.rowptr[i + 1];
        double d = 0.0;
        for (int k = ks; k < ke; ++k) if (A.col[k] == i) d = __dadd_rn(d, A.val[k]);
        double l = d;
        int o = f_rowptr[i] + 1;
        for (int k = ks; k < ke; ++k) {
            const int j = A.col[k];
            if (j == i) continue;
            const double a = A.val[k];
            if (is_strong(a, rmi, rowmax[j], theta)) { f_val[o] = a; f_col[o] = j; ++o; }
            else l = __dadd_rn(l, a);
        }
        double dF = (l > 0.0) ? l : d;
        if (dF == 0.0) dF = 1.0;
        f_val[f_rowptr[i]] = dF; f_col[f_rowptr[i]] = i;
        dinvF[i] = 1.0 / dF;
    }
}

// A -> filtered matrix F (owned) and 1 / its diagonal
int strength_filter(mgcg_ctx *c, const Csr &A, double theta, DevCsr *F, double **dinvF)
{
    cudaStream_t st = c->stream;
    const int n = A.rows, g = agrid(n + 1, c->sm_count);
    double *rowmax = nullptr;
    int *cnt = nullptr;
    MGCG_CUDA(pool_alloc((void **)&rowmax, sizeof(double) * (size_t)n, st));
    MGCG_CUDA(pool_alloc((void **)&cnt, sizeof(int) * ((size_t)n + 1), st));
    k_rowmax<<<g, 256, 0, st>>>(A, rowmax);
    k_filter_count<<<g, 256, 0, st>>>(A, rowmax, theta, cnt);
    MGCG_LAUNCH_CHECK();
    F->rows = n; F->cols = n;
    MGCG_CUDA(pool_alloc((void **)&F->rowptr, sizeof(int) * ((size_t)n + 1), st));
    int nnz = 0;
    {
        const int s_ = scan_counts_checked(c, cnt, n, F->rowptr, &nnz, "strength filter");
        pool_free(cnt, st);
        if (s_ != MGCG_OK) { pool_free(rowmax, st); return s_; }
    }
    F->nnz = nnz;
    MGCG_CUDA(pool_alloc((void **)&F->val, sizeof(double) * (size_t)(nnz > 0 ? nnz : 1), st));
    MGCG_CUDA(pool_alloc((void **)&F->col, sizeof(int) * (size_t)(nnz > 0 ? nnz : 1), st));
    MGCG_CUDA(pool_alloc((void **)dinvF, sizeof(double) * (size_t)n, st));
    k_filter_fill<<<g, 256, 0, st>>>(A, rowmax, theta, F->rowptr, F->val, F->col, *dinvF);
    MGCG_LAUNCH_CHECK();
    pool_free(rowmax, st);
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_mis2_pack(int n, const unsigned char *__restrict__ state, u64 *__restrict__ T0, int *undecided)
{
    int mine = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned s = state[i];
        T0[i] = ((u64)s << 62) | ((u64)(mix32((unsigned)i) >> 2) << 32) | (u64)(unsigned)i;
        mine += (s == 1u);
    }
    for (int o = 16; o > 0; o >>= 1) mine += __shfl_down_sync(0xffffffffu, mine, o);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(undecided, mine);
}

// Tout_i = max(Tin_i, max over the strong neighbours of i) — F stores the diagonal first, the neighbours after it
__global__ void __launch_bounds__(256) k_mis2_hop(Csr F, const u64 *__restrict__ Tin, u64 *__restrict__ Tout)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < F.rows; i += gridDim.x * blockDim.x) {
        u64 m = Tin[i];
        for (int k = F.rowptr[i] + 1; k < F.rowptr[i + 1]; ++k) { const u64 t = Tin[F.col[k]]; if (t > m) m = t; }
        Tout[i] = m;
    }
}

__global__ void __launch_bounds__(256) k_mis2_update(int n, unsigned char *__restrict__ state, const u64 *__restrict__ T0, const u64 *__restrict__ T2)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        if (state[i] != 1u) continue;
        const u64 t2 = T2[i];
        if (t2 == T0[i]) state[i] = 2u;
        else if ((t2 >> 62) == 2ull) state[i] = 0u;
    }
}

__global__ void __launch_bounds__(256) k_flag_state(int n, const unsigned char *__restrict__ state, unsigned char want, int *__restrict__ flag)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= n; i += gridDim.x * blockDim.x) flag[i] = (i < n && state[i] == want) ? 1 : 0;
}
__global__ void __launch_bounds__(256) k_flag_negative(int n, const int *__restrict__ agg, int *__restrict__ flag)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= n; i += gridDim.x * blockDim.x) flag[i] = (i < n && agg[i] < 0) ? 1 : 0;
}
__global__ void __launch_bounds__(256) k_agg_roots(int n, const unsigned char *__restrict__ state, const int *__restrict__ rid, int *__restrict__ agg)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) agg[i] = (state[i] == 2u) ? rid[i] : -1;
}
// dst_i = src_i if aggregated, else the aggregate of the strong neighbour with the largest |a_ij| among the aggregated ones
__global__ void __launch_bounds__(256) k_agg_attach(Csr F, const int *__restrict__ src, int *__restrict__ dst)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < F.rows; i += gridDim.x * blockDim.x) {
        int pick = src[i];
        if (pick < 0) {
            double best = -1.0;
            for (int k = F.rowptr[i] + 1; k < F.rowptr[i + 1]; ++k) {
                const int a_j = src[F.col[k]];
                if (a_j < 0) continue;
                const double a = fabs(F.val[k]);
                if (a > best) { best = a; pick = a_j; }
            }
        }
        dst[i] = pick;
    }
}
__global__ void __launch_bounds__(256) k_agg_leftover(int n, int base, const int *__restrict__ rank, int *__restrict__ agg)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        if (agg[i] < 0) agg[i] = base + rank[i];
}

static int exclusive_scan_int(mgcg_ctx *c, const int *in, int *out, int count)
{
    cudaStream_t st = c->stream;
    size_t tb = 0;
    MGCG_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, in, out, count, st));
    void *tmp = nullptr;
    MGCG_CUDA(pool_alloc(&tmp, tb, st));
    MGCG_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tb, in, out, count, st));
    pool_free(tmp, st);
    return MGCG_OK;
}

// F: filtered matrix (diagonal first).  agg[0..n) <- aggregate ids, *ncoarse <- their number.
int mis2_aggregate(mgcg_ctx *c, const Csr &F, int *agg, int *ncoarse, int *rounds_out)
{
    cudaStream_t st = c->stream;
    const int n = F.rows, g = agrid(n + 1, c->sm_count);
    unsigned char *state = nullptr;
    u64 *T0 = nullptr, *T1 = nullptr, *T2 = nullptr;
    int *d_cnt = nullptr, *flag = nullptr, *rank = nullptr, *agg2 = nullptr;
    MGCG_CUDA(pool_alloc((void **)&state, (size_t)n, st));
    MGCG_CUDA(pool_alloc((void **)&T0, sizeof(u64) * (size_t)n, st));
    MGCG_CUDA(pool_alloc((void **)&T1, sizeof(u64) * (size_t)n, st));
    MGCG_CUDA(pool_alloc((void **)&T2, sizeof(u64) * (size_t)n, st));
    MGCG_CUDA(pool_alloc((void **)&d_cnt, sizeof(int), st));
    MGCG_CUDA(pool_alloc((void **)&flag, sizeof(int) * ((size_t)n + 1), st));
    MGCG_CUDA(pool_alloc((void **)&rank, sizeof(int) * ((size_t)n + 1), st));
    MGCG_CUDA(pool_alloc((void **)&agg2, sizeof(int) * (size_t)n, st));
    MGCG_CUDA(cudaMemsetAsync(state, 1, (size_t)n, st));
    int rounds = 0;
    for (; rounds < 256; ++rounds) {
        int undecided = 0;
        MGCG_CUDA(cudaMemsetAsync(d_cnt, 0, sizeof(int), st));
        k_mis2_pack<<<g, 256, 0, st>>>(n, state, T0, d_cnt);
        MGCG_LAUNCH_CHECK();
        MGCG_CUDA(cudaMemcpyAsync(&undecided, d_cnt, sizeof(int), cudaMemcpyDeviceToHost, st));
        MGCG_CUDA(cudaStreamSynchronize(st));
        if (!undecided) break;
        k_mis2_hop<<<g, 256, 0, st>>>(F, T0, T1);
        k_mis2_hop<<<g, 256, 0, st>>>(F, T1, T2);
        k_mis2_update<<<g, 256, 0, st>>>(n, state, T0, T2);
        MGCG_LAUNCH_CHECK();
    }
    if (rounds_out) *rounds_out = rounds;
    // roots in index order
    k_flag_state<<<g, 256, 0, st>>>(n, state, 2u, flag);
    MGCG_LAUNCH_CHECK();
    MGCG_CHECK(exclusive_scan_int(c, flag, rank, n + 1));
    int nroots = 0, nleft = 0;
    MGCG_CUDA(cudaMemcpyAsync(&nroots, rank + n, sizeof(int), cudaMemcpyDeviceToHost, st));
    k_agg_roots<<<g, 256, 0, st>>>(n, state, rank, agg);
    k_agg_attach<<<g, 256, 0, st>>>(F, agg, agg2);
    k_agg_attach<<<g, 256, 0, st>>>(F, agg2, agg);
    // leftovers (no aggregated strong neighbour after two passes): singletons numbered after the roots, in index order
    k_flag_negative<<<g, 256, 0, st>>>(n, agg, flag);
    MGCG_LAUNCH_CHECK();
    MGCG_CHECK(exclusive_scan_int(c, flag, rank, n + 1));
    MGCG_CUDA(cudaMemcpyAsync(&nleft, rank + n, sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    if (nleft > 0) {
        k_agg_leftover<<<g, 256, 0, st>>>(n, nroots, rank, agg);
        MGCG_LAUNCH_CHECK();
    }
    *ncoarse = nroots + nleft;
    pool_free(state, st); pool_free(T0, st); pool_free(T1, st); pool_free(T2, st);
    pool_free(d_cnt, st); pool_free(flag, st); pool_free(rank, st); pool_free(agg2, st);
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
// red-black Gauss-Seidel needs a two-colouring: with rows = cells of a grid, parity (ix + iy + iz) & 1 is one exactly when no
// off-diagonal entry joins two cells of the same parity (5-/7-point stencils and their plain-aggregation Galerkin products; not
// the 27-point operator, not a smoothed-aggregation coarse level).
__global__ void __launch_bounds__(256) k_check_two_colour(Csr A, int nx, int ny, int *bad)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.rows; i += gridDim.x * blockDim.x) {
        const int pi = ((i % nx) + ((i / nx) % ny) + (i / nx / ny)) & 1;
        for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k) {
            const int j = A.col[k];
            if (j == i) continue;
            const int pj = ((j % nx) + ((j / nx) % ny) + (j / nx / ny)) & 1;
            if (pj == pi) { *bad = 1; break; }
        }
    }
}

int check_two_colourable(mgcg_ctx *c, const Csr &A, int nx, int ny, int *result)
{
    cudaStream_t st = c->stream;
    *result = 0;
    if (nx <= 0 || ny <= 0) return MGCG_OK;
    int *d_bad = nullptr, bad = 0;
    MGCG_CUDA(pool_alloc((void **)&d_bad, sizeof(int), st));
    MGCG_CUDA(cudaMemsetAsync(d_bad, 0, sizeof(int), st));
    k_check_two_colour<<<agrid(A.rows, c->sm_count), 256, 0, st>>>(A, nx, ny, d_bad);
    MGCG_LAUNCH_CHECK();
    MGCG_CUDA(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    pool_free(d_bad, st);
    *result = bad ? 0 : 1;
    return MGCG_OK;
}

}  // namespace mgcg
