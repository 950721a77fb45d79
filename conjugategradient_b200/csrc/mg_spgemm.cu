// mg_spgemm.cu — row-wise sparse matrix product C = A * B for the hierarchy build (Galerkin products A*P and R*(A*P)).
// (No counterpart in the reference; CPU twin: oracle/mg_ref.c spgemm.)
//
// A group of G lanes owns one row of C.  It walks the row of A entry by entry IN STORAGE ORDER; for each a_ik the lanes take
// the entries of row k of B (distinct columns) and add a_ik*b_kj into a per-group hash table in shared memory.  Because
// the walk over k is sequential and the lanes of one step hit distinct columns, every c_ij is the sum of its products in
// storage order of the A row, each product rounded before the add (__dmul_rn/__dadd_rn, no FMA) — the same order and the
// same roundings as the CPU twin, so the coarse operators are bit-identical and run-to-run deterministic without a global
// sort.  Two phases: symbolic (unique columns per row) -> scan -> numeric (fill, bitonic-sort the row by column, write).
// Rows whose product count exceeds the largest table fall back to the expand/sort/compress path (spgemm_esc).
#include "mg.cuh"
#include <cub/cub.cuh>
#include <stdlib.h>

namespace mgcg {

constexpr int kEmpty = -1;

__device__ __forceinline__ unsigned hash_col(int col, int cap_mask) { return ((unsigned)col * 0x9E3779B1u >> 7) & (unsigned)cap_mask; }

template <int G>
__device__ __forceinline__ unsigned group_mask()
{
    if (G == 32) return 0xffffffffu;
    const int lane = threadIdx.x & 31;
    return ((1u << G) - 1u) << (lane & ~(G - 1));
}

// One pass over row `row` of A*B for a group of G lanes.  The lanes first load up to G entries of the A row together with
// the extents of the B rows they select (one memory latency for G entries instead of a dependent chain per entry), then
// walk them in storage order by shuffle broadcast; the first G entries of the NEXT B row are prefetched while the current
// ones are inserted.  NUMERIC = false: count distinct columns (returns this lane's share);  true: accumulate values.
template <int G, bool NUMERIC>
__device__ __forceinline__ int spgemm_row(const Csr &A, const Csr &B, int shift, int row, int lane, unsigned mask, int cap_mask,
                                          int *keys, double *vals, int *overflow)
{
    int mine = 0;
    const int ks = A.rowptr[row], ke = A.rowptr[row + 1];
    for (int kb = ks; kb < ke; kb += G) {
        const int kk = kb + lane;
        int q0_l = 0, q1_l = 0;
        double a_l = 0.0;
        if (kk < ke) {
            const int j = A.col[kk] - shift;
            if (NUMERIC) a_l = A.val[kk];
            q0_l = B.rowptr[j]; q1_l = B.rowptr[j + 1];
        }
        const int nk = min(G, ke - kb);
        int q0 = __shfl_sync(mask, q0_l, 0, G), q1 = __shfl_sync(mask, q1_l, 0, G);
        int c_cur = 0; double v_cur = 0.0;
        if (q0 + lane < q1) { c_cur = B.col[q0 + lane]; if (NUMERIC) v_cur = B.val[q0 + lane]; }
        for (int t = 0; t < nk; ++t) {
            const double a = NUMERIC ? __shfl_sync(mask, a_l, t, G) : 0.0;
            int q0n = 0, q1n = 0, c_nxt = 0; double v_nxt = 0.0;
            if (t + 1 < nk) {
                q0n = __shfl_sync(mask, q0_l, t + 1, G); q1n = __shfl_sync(mask, q1_l, t + 1, G);
                if (q0n + lane < q1n) { c_nxt = B.col[q0n + lane]; if (NUMERIC) v_nxt = B.val[q0n + lane]; }
            }
            for (int q = q0 + lane; q < q1; q += G) {
                int col = c_cur; double bv = v_cur;
                if (q != q0 + lane) { col = B.col[q]; if (NUMERIC) bv = B.val[q]; }   // B rows longer than G (rare)
                unsigned s = hash_col(col, cap_mask);
                int probes = 0;
                while (true) {
                    const int old = atomicCAS(&keys[s], kEmpty, col);
                    if (old == kEmpty) { ++mine; break; }
                    if (old == col) break;
                    s = (s + 1) & cap_mask;
                    if (!NUMERIC && ++probes > 48) { *overflow = 1; break; }   // table (nearly) full: the host retries larger
                }
                if (NUMERIC) vals[s] = __dadd_rn(vals[s], __dmul_rn(a, bv));   // first product: 0 + prod == prod exactly
            }
            __syncwarp(mask);   // orders this step's table writes before the next step's reads
            q0 = q0n; q1 = q1n; c_cur = c_nxt; v_cur = v_nxt;
        }
    }
    return mine;
}

// ---- symbolic: cnt[row] = number of distinct columns of row `row` of A*B
template <int G>
__global__ void __launch_bounds__(256) k_spgemm_symbolic(Csr A, Csr B, int shift, int cap, int *__restrict__ cnt, int *__restrict__ overflow)
{
    extern __shared__ int sm_keys[];
    const int gpb = blockDim.x / G;
    const int lane = threadIdx.x & (G - 1);
    int *keys = sm_keys + (size_t)(threadIdx.x / G) * (cap + 1);   // +1: the groups of a warp start in different banks
    const unsigned mask = group_mask<G>();
    for (int row = blockIdx.x * gpb + threadIdx.x / G; row < A.rows; row += gridDim.x * gpb) {
        for (int s = lane; s < cap; s += G) keys[s] = kEmpty;
        __syncwarp(mask);
        // counting needs no summation order: every lane takes one entry of the A row and walks the B row it selects, so all
        // lanes insert concurrently (the numeric phase below must go entry by entry instead)
        int mine = 0;
        const int cap_mask = cap - 1;
        const int ks = A.rowptr[row], ke = A.rowptr[row + 1];
        for (int kk = ks + lane; kk < ke; kk += G) {
            const int j = A.col[kk] - shift;
            const int q1 = B.rowptr[j + 1];
            for (int q = B.rowptr[j]; q < q1; ++q) {
                const int col = B.col[q];
                unsigned s = hash_col(col, cap_mask);
                int probes = 0;
                while (true) {
                    const int old = atomicCAS(&keys[s], kEmpty, col);
                    if (old == kEmpty) { ++mine; break; }
                    if (old == col) break;
                    s = (s + 1) & cap_mask;
                    if (++probes > 48) { *overflow = 1; break; }   // table (nearly) full: the host retries with a larger one
                }
            }
        }
        __syncwarp(mask);
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) mine += __shfl_down_sync(mask, mine, o, G);
        if (lane == 0) cnt[row] = mine;
        __syncwarp(mask);
    }
}

// ---- numeric: fill row `row` of C (columns ascending)
template <int G>
__global__ void __launch_bounds__(256) k_spgemm_numeric(Csr A, Csr B, int shift, int cap, const int *__restrict__ c_rowptr,
                                                        int *__restrict__ c_col, double *__restrict__ c_val)
{
    extern __shared__ double sm_raw[];
    const int gpb = blockDim.x / G;
    const int lane = threadIdx.x & (G - 1);
    double *vals = sm_raw + (size_t)(threadIdx.x / G) * (cap + 1);   // +1: bank-conflict padding between groups
    int *keys = reinterpret_cast<int *>(sm_raw + (size_t)gpb * (cap + 1)) + (size_t)(threadIdx.x / G) * (cap + 1);
    const unsigned mask = group_mask<G>();
    for (int row = blockIdx.x * gpb + threadIdx.x / G; row < A.rows; row += gridDim.x * gpb) {
        for (int s = lane; s < cap; s += G) { keys[s] = kEmpty; vals[s] = 0.0; }
        __syncwarp(mask);
        spgemm_row<G, true>(A, B, shift, row, lane, mask, cap - 1, keys, vals, nullptr);
        // empty slots sort to the end
        for (int s = lane; s < cap; s += G) if (keys[s] == kEmpty) keys[s] = 0x7fffffff;
        __syncwarp(mask);
        for (int kk = 2; kk <= cap; kk <<= 1) {
            for (int jj = kk >> 1; jj > 0; jj >>= 1) {
                for (int pr = lane; pr < (cap >> 1); pr += G) {     // one compare-exchange per pair
                    const int idx = ((pr & ~(jj - 1)) << 1) | (pr & (jj - 1)), partner = idx | jj;
                    const int ka = keys[idx], kb = keys[partner];
                    const bool up = (idx & kk) == 0;
                    if ((ka > kb) == up) {
                        keys[idx] = kb; keys[partner] = ka;
                        const double va = vals[idx]; vals[idx] = vals[partner]; vals[partner] = va;
                    }
                }
                __syncwarp(mask);
            }
        }
        const int o0 = c_rowptr[row], n_out = c_rowptr[row + 1] - o0;
        for (int s = lane; s < n_out; s += G) { c_col[o0 + s] = keys[s]; c_val[o0 + s] = vals[s]; }
        __syncwarp(mask);
    }
}

// max over rows of the product count (upper bound of the distinct columns of a row of C), clipped to int
__global__ void k_spgemm_maxprod(Csr A, Csr B, int shift, int *out_max)
{
    int m = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.rows; i += gridDim.x * blockDim.x) {
        long long s = 0;
        for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k) { const int j = A.col[k] - shift; s += B.rowptr[j + 1] - B.rowptr[j]; }
        m = max(m, (int)min(s, (long long)0x7fffffff));
    }
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_down_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(out_max, m);
}

// ---------------------------------------------------------------------------------------------------------
// Row counts -> row pointers with the total accumulated in 64 bits.  A*P and R*(A*P) hold 2-3x nnz(A) entries, so a
// problem that itself fits int32 can produce a product that does not; the int32 exclusive scan would wrap silently and
// size C->val / C->col from the wrapped total.  The total is summed separately as unsigned long long and any level with
// >= 2^31 entries is refused (MGCG_ERR_UNSUPPORTED) before anything is allocated from it.
__global__ void __launch_bounds__(256) k_sum_counts64(const int *__restrict__ cnt, int n, unsigned long long *out)
{
    unsigned long long s = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) s += (unsigned long long)(unsigned)cnt[i];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0 && s) atomicAdd(out, s);
}

// cnt[0..rows) -> rowptr[0..rows] (cnt[rows] must be 0); *nnz = total.  Synchronises the stream.
int scan_counts_checked(mgcg_ctx *c, const int *cnt, int rows, int *rowptr, int *nnz, const char *what)
{
    cudaStream_t st = c->stream;
    unsigned long long *d_total = nullptr, h_total = 0;
    MGCG_CUDA(pool_alloc((void **)&d_total, sizeof(unsigned long long), st));
    MGCG_CUDA(cudaMemsetAsync(d_total, 0, sizeof(unsigned long long), st));
    int g = (rows + 255) / 256;
    if (g > c->sm_count * 8) g = c->sm_count * 8;
    if (g < 1) g = 1;
    k_sum_counts64<<<g, 256, 0, st>>>(cnt, rows, d_total);
    MGCG_LAUNCH_CHECK();
    size_t tb = 0;
    MGCG_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt, rowptr, rows + 1, st));
    void *tmp = nullptr;
    MGCG_CUDA(pool_alloc(&tmp, tb, st));
    MGCG_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tb, cnt, rowptr, rows + 1, st));
    MGCG_CUDA(cudaMemcpyAsync(&h_total, d_total, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    pool_free(tmp, st); pool_free(d_total, st);
    if (h_total >= (1ull << 31)) {
        set_error("%s: %llu entries exceed the int32 index range of a level operator (2^31 - 1)", what, h_total);
        return MGCG_ERR_UNSUPPORTED;
    }
    *nnz = (int)h_total;
    return MGCG_OK;
}


// ---------------------------------------------------------------------------------------------------------
// FEW, VERY LONG rows (the coarsest levels: a few hundred to a few thousand rows whose products run to 10^5 per row, all
// columns within a few thousand): one BLOCK per row with a dense accumulator over the column space in shared memory.
// The A row is still walked entry by entry in storage order (a block barrier per entry orders the adds to one c_ij), the
// 256 threads take the entries of the selected B row (distinct columns, no conflicts), so the sums and roundings are the
// ones of the hash kernels and of the CPU twin.  No hashing, no sort: the output is read off the accumulator in column
// order.  With 32 lanes per row these levels took 3-6 ms per product (9 300 dependent hash steps per row at 216 rows).
constexpr int kDenseThreads = 256;
constexpr int kDenseChunk = 256;      // A-row entries staged per pass

template <bool NUMERIC>
__global__ void __launch_bounds__(kDenseThreads) k_spgemm_dense_row(Csr A, Csr B, int shift, int cols, int *__restrict__ cnt,
                                                                    const int *__restrict__ c_rowptr, int *__restrict__ c_col,
                                                                    double *__restrict__ c_val)
{
    extern __shared__ double sm_dense[];
    double *acc = sm_dense;                                             // [cols]              (NUMERIC)
    unsigned char *mark = reinterpret_cast<unsigned char *>(sm_dense + (NUMERIC ? cols : 0));   // [cols]
    __shared__ int s_q0[kDenseChunk], s_q1[kDenseChunk];
    __shared__ double s_a[kDenseChunk];
    __shared__ int s_scan[kDenseThreads];
    const int t = threadIdx.x;
    for (int c = t; c < cols; c += kDenseThreads) { if (NUMERIC) acc[c] = 0.0; mark[c] = 0; }
    __syncthreads();
    for (int row = blockIdx.x; row < A.rows; row += gridDim.x) {
        const int ks = A.rowptr[row], ke = A.rowptr[row + 1];
        for (int kb = ks; kb < ke; kb += kDenseChunk) {
            const int nk = min(kDenseChunk, ke - kb);
            if (t < nk) {
                const int j = A.col[kb + t] - shift;
                s_q0[t] = B.rowptr[j]; s_q1[t] = B.rowptr[j + 1];
                if (NUMERIC) s_a[t] = A.val[kb + t];
            }
            __syncthreads();
            if (!NUMERIC) {   // counting needs no order: every thread takes whole B rows
                for (int e = t; e < nk; e += kDenseThreads)
                    for (int q = s_q0[e]; q < s_q1[e]; ++q) mark[B.col[q]] = 1;
            } else {
                int q0 = s_q0[0], q1 = s_q1[0];
                int c_cur = 0; double v_cur = 0.0;
                if (q0 + t < q1) { c_cur = B.col[q0 + t]; v_cur = B.val[q0 + t]; }
                for (int e = 0; e < nk; ++e) {
                    const double a = s_a[e];
                    int q0n = 0, q1n = 0, c_nxt = 0; double v_nxt = 0.0;
                    if (e + 1 < nk) {   // the next B row's first entries are in flight while this one is added
                        q0n = s_q0[e + 1]; q1n = s_q1[e + 1];
                        if (q0n + t < q1n) { c_nxt = B.col[q0n + t]; v_nxt = B.val[q0n + t]; }
                    }
                    for (int q = q0 + t; q < q1; q += kDenseThreads) {
                        int col = c_cur; double bv = v_cur;
                        if (q != q0 + t) { col = B.col[q]; bv = B.val[q]; }
                        acc[col] = __dadd_rn(acc[col], __dmul_rn(a, bv));      // first product: 0 + prod == prod exactly
                        mark[col] = 1;
                    }
                    __syncthreads();      // orders this entry's adds before the next entry's
                    q0 = q0n; q1 = q1n; c_cur = c_nxt; v_cur = v_nxt;
                }
            }
            __syncthreads();
        }
        // ---- read the row off in column order: thread t owns the contiguous slice [c0, c1)
        const int per = (cols + kDenseThreads - 1) / kDenseThreads;
        const int c0 = min(t * per, cols), c1 = min(c0 + per, cols);
        int mine = 0;
        for (int c = c0; c < c1; ++c) mine += mark[c];
        s_scan[t] = mine;
        __syncthreads();
        for (int o = 1; o < kDenseThreads; o <<= 1) {      // inclusive Hillis-Steele scan
            const int v = (t >= o) ? s_scan[t - o] : 0;
            __syncthreads();
            s_scan[t] += v;
            __syncthreads();
        }
        if (!NUMERIC) {
            if (t == kDenseThreads - 1) cnt[row] = s_scan[t];
            for (int c = c0; c < c1; ++c) mark[c] = 0;
        } else {
            int o = c_rowptr[row] + s_scan[t] - mine;
            for (int c = c0; c < c1; ++c)
                if (mark[c]) { c_col[o] = c; c_val[o] = acc[c]; ++o; acc[c] = 0.0; mark[c] = 0; }
        }
        __syncthreads();
    }
}

// C = A*B for few long rows over a small column space.  *done = false when the shape does not qualify.
static int run_dense_rows(mgcg_ctx *c, const Csr &A, const Csr &B, int shift, int maxprod, DevCsr *C, bool *done)
{
    cudaStream_t st = c->stream;
    *done = false;
    const int cols = B.cols;
    int min_prod = 16384;      // below this the 32-lane hash kernels are faster (MGCG_SPGEMM_DENSE_MIN: tests force the path)
    if (const char *e = getenv("MGCG_SPGEMM_DENSE_MIN")) min_prod = atoi(e);
    if (A.rows > 2048 || cols > 8192 || maxprod < min_prod) return MGCG_OK;
    const size_t sm_s = (size_t)cols + 16, sm_n = (size_t)cols * 9 + 16;
    MGCG_CUDA(cudaFuncSetAttribute(k_spgemm_dense_row<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_n));
    int *cnt = nullptr;
    MGCG_CUDA(pool_alloc((void **)&cnt, sizeof(int) * ((size_t)A.rows + 1), st));
    MGCG_CUDA(cudaMemsetAsync(cnt + A.rows, 0, sizeof(int), st));
    int grid = A.rows < c->sm_count * 2 ? A.rows : c->sm_count * 2;
    k_spgemm_dense_row<false><<<grid, kDenseThreads, sm_s, st>>>(A, B, shift, cols, cnt, nullptr, nullptr, nullptr);
    MGCG_LAUNCH_CHECK();
    C->rows = A.rows; C->cols = B.cols;
    MGCG_CUDA(pool_alloc((void **)&C->rowptr, sizeof(int) * ((size_t)A.rows + 1), st));
    int nnz = 0;
    {
        const int s_ = scan_counts_checked(c, cnt, A.rows, C->rowptr, &nnz, "spgemm");
        pool_free(cnt, st);
        if (s_ != MGCG_OK) { pool_free(C->rowptr, st); C->rowptr = nullptr; return s_; }
    }
    C->nnz = nnz;
    MGCG_CUDA(pool_alloc((void **)&C->val, sizeof(double) * (size_t)(nnz > 0 ? nnz : 1), st));
    MGCG_CUDA(pool_alloc((void **)&C->col, sizeof(int) * (size_t)(nnz > 0 ? nnz : 1), st));
    if (getenv("MGCG_TRACE")) fprintf(stderr, "[mgcg spgemm] rows %d dense-row kernel, cols %d maxprod %d nnz %d grid %d\n", A.rows, cols, maxprod, nnz, grid);
    k_spgemm_dense_row<true><<<grid, kDenseThreads, sm_n, st>>>(A, B, shift, cols, nullptr, C->rowptr, C->col, C->val);
    MGCG_LAUNCH_CHECK();
    *done = true;
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
// SHORT A rows (<= MAXA entries, the stencil levels): one THREAD per row does a k-way merge of the B rows its A entries
// select.  B rows are sorted by column (every producer in this library emits them so), so the smallest head column is the
// next output column, and adding the heads that carry it IN k ORDER is exactly the storage-order sum of the hash kernels
// and of the CPU twin.  No table, no sort, no shared memory: two streaming passes (count, fill).  At 256^3 the fine-level
// A*P (16.7 M rows, 105 M outputs) took 3.7 + 13.0 ms with the 4-lane hash kernels.
template <int MAXA, bool NUMERIC>
__global__ void __launch_bounds__(128) k_spgemm_merge(Csr A, Csr B, int shift, int *__restrict__ cnt, const int *__restrict__ c_rowptr,
                                                      int *__restrict__ c_col, double *__restrict__ c_val)
{
    constexpr int kEnd = 0x7fffffff;
    for (int row = blockIdx.x * blockDim.x + threadIdx.x; row < A.rows; row += gridDim.x * blockDim.x) {
        int q[MAXA], qe[MAXA], hc[MAXA];
        double a[MAXA];
        const int ks = A.rowptr[row], m = A.rowptr[row + 1] - ks;
#pragma unroll
        for (int k = 0; k < MAXA; ++k) {
            q[k] = 0; qe[k] = 0; hc[k] = kEnd; a[k] = 0.0;
            if (k < m) {
                const int j = A.col[ks + k] - shift;
                q[k] = B.rowptr[j]; qe[k] = B.rowptr[j + 1];
                if (NUMERIC) a[k] = A.val[ks + k];
                if (q[k] < qe[k]) hc[k] = B.col[q[k]];
            }
        }
        int o = NUMERIC ? c_rowptr[row] : 0;
        while (true) {
            int cur = kEnd;
#pragma unroll
            for (int k = 0; k < MAXA; ++k) cur = min(cur, hc[k]);
            if (cur == kEnd) break;
            double s = 0.0;
            bool have = false;
#pragma unroll
            for (int k = 0; k < MAXA; ++k)
                if (hc[k] == cur) {
                    if (NUMERIC) {
                        const double pr = __dmul_rn(a[k], B.val[q[k]]);
                        s = have ? __dadd_rn(s, pr) : pr;
                        have = true;
                    }
                    ++q[k];
                    hc[k] = q[k] < qe[k] ? B.col[q[k]] : kEnd;
                }
            if (NUMERIC) { c_col[o] = cur; c_val[o] = s; }
            ++o;
        }
        if (!NUMERIC) cnt[row] = o;
    }
}

__global__ void k_spgemm_maxlen(Csr A, int *out_max)
{
    int m = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.rows; i += gridDim.x * blockDim.x) m = max(m, A.rowptr[i + 1] - A.rowptr[i]);
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_down_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(out_max, m);
}

template <int MAXA>
static int run_merge(mgcg_ctx *c, const Csr &A, const Csr &B, int shift, DevCsr *C)
{
    cudaStream_t st = c->stream;
    int *cnt = nullptr;
    MGCG_CUDA(pool_alloc((void **)&cnt, sizeof(int) * ((size_t)A.rows + 1), st));
    MGCG_CUDA(cudaMemsetAsync(cnt + A.rows, 0, sizeof(int), st));
    int grid = (A.rows + 127) / 128;
    if (grid > c->sm_count * 64) grid = c->sm_count * 64;
    k_spgemm_merge<MAXA, false><<<grid, 128, 0, st>>>(A, B, shift, cnt, nullptr, nullptr, nullptr);
    MGCG_LAUNCH_CHECK();
    C->rows = A.rows; C->cols = B.cols;
    MGCG_CUDA(pool_alloc((void **)&C->rowptr, sizeof(int) * ((size_t)A.rows + 1), st));
    int nnz = 0;
    {
        const int s_ = scan_counts_checked(c, cnt, A.rows, C->rowptr, &nnz, "spgemm");
        pool_free(cnt, st);
        if (s_ != MGCG_OK) { pool_free(C->rowptr, st); C->rowptr = nullptr; return s_; }
    }
    C->nnz = nnz;
    MGCG_CUDA(pool_alloc((void **)&C->val, sizeof(double) * (size_t)(nnz > 0 ? nnz : 1), st));
    MGCG_CUDA(pool_alloc((void **)&C->col, sizeof(int) * (size_t)(nnz > 0 ? nnz : 1), st));
    if (getenv("MGCG_TRACE")) fprintf(stderr, "[mgcg spgemm] rows %d merge kernel MAXA %d nnz %d grid %d\n", A.rows, MAXA, nnz, grid);
    k_spgemm_merge<MAXA, true><<<grid, 128, 0, st>>>(A, B, shift, nullptr, C->rowptr, C->col, C->val);
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
// MEDIUM A rows (33..128 entries) over SHORT B rows (the fine-level Galerkin product R*(A*P): 636 k rows of ~79 entries
// selecting AP rows of ~6): one WARP per row does the same k-way merge, each lane holding the heads of 4 consecutive
// lists.  The next output column is a warp-wide minimum (redux.sync); the lanes whose heads carry it drop their products
// into a per-warp shared-memory slot indexed by list number, and the sum is taken over the set bits of the four ballots in
// list order — storage order of the A row, product by product, like the hash kernels and the CPU twin.
// At 256^3 this product took 2.5 + 6.9 ms with the 8-lane hash kernels.
constexpr int kWmSlots = 4;       // 33..128 entries per A row
constexpr int kWmSlotsLong = 8;   // 129..256 entries per A row (the reference's 160-per-row band, particle matrices at r_e = 3.1 l0)
constexpr int kWmWarps = 8;

template <bool NUMERIC, int SLOTS>
__global__ void __launch_bounds__(kWmWarps * 32) k_spgemm_wmerge(Csr A, Csr B, int shift, int *__restrict__ cnt, const int *__restrict__ c_rowptr,
                                                                int *__restrict__ c_col, double *__restrict__ c_val)
{
    constexpr unsigned kEnd = 0x7fffffffu;
    __shared__ double s_prod[kWmWarps][32 * SLOTS];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int nwarps = gridDim.x * kWmWarps;
    for (int row = blockIdx.x * kWmWarps + w; row < A.rows; row += nwarps) {
        int q[SLOTS], qe[SLOTS];
        unsigned hc[SLOTS];
        double a[SLOTS];
        const int ks = A.rowptr[row], m = A.rowptr[row + 1] - ks;
#pragma unroll
        for (int t = 0; t < SLOTS; ++t) {
            const int k = lane * SLOTS + t;
            q[t] = 0; qe[t] = 0; hc[t] = kEnd; a[t] = 0.0;
            if (k < m) {
                const int j = A.col[ks + k] - shift;
                q[t] = B.rowptr[j]; qe[t] = B.rowptr[j + 1];
                if (NUMERIC) a[t] = A.val[ks + k];
                if (q[t] < qe[t]) hc[t] = (unsigned)B.col[q[t]];
            }
        }
        int o = NUMERIC ? c_rowptr[row] : 0;
        while (true) {
            unsigned mine = kEnd;
#pragma unroll
            for (int t = 0; t < SLOTS; ++t) mine = min(mine, hc[t]);
            const unsigned cur = __reduce_min_sync(0xffffffffu, mine);
            if (cur == kEnd) break;
            unsigned bal[SLOTS];
            unsigned any = 0u;
#pragma unroll
            for (int t = 0; t < SLOTS; ++t) {
                const bool hit = hc[t] == cur;
                if (NUMERIC && hit) s_prod[w][lane * SLOTS + t] = __dmul_rn(a[t], B.val[q[t]]);
                bal[t] = __ballot_sync(0xffffffffu, hit);
                any |= bal[t];
                if (hit) { ++q[t]; hc[t] = q[t] < qe[t] ? (unsigned)B.col[q[t]] : kEnd; }
            }
            if (NUMERIC) {
                __syncwarp();
                double sum = 0.0;
                bool have = false;
                while (any) {          // lanes in ascending order, slots in ascending order == list number == storage order of the A row
                    const int l = __ffs(any) - 1;
                    any &= any - 1;
#pragma unroll
                    for (int t = 0; t < SLOTS; ++t)
                        if ((bal[t] >> l) & 1u) {
                            const double v = s_prod[w][l * SLOTS + t];
                            sum = have ? __dadd_rn(sum, v) : v;
                            have = true;
                        }
                }
                if (lane == 0) { c_col[o] = (int)cur; c_val[o] = sum; }
                __syncwarp();      // the slots are rewritten by the next output
            }
            ++o;
        }
        if (!NUMERIC && lane == 0) cnt[row] = o;
    }
}


// Numeric pass for output rows of at most 32 entries (the fine-level Galerkin product: 27 per row): OWNER lanes.
// Phase 1 repeats the symbolic merge and leaves the row's sorted distinct columns one per lane.  Phase 2 walks the A row in storage
// order, 32 entries at a time; every lane reads the B row of the current entry (uniform addresses: one L1 transaction per load) and
// the lane that owns the entry's column adds the product — so c_ij is still summed in the storage order of the A row, product by
// product, bit-identical to the merge kernels and the CPU twin, but the ordered sums of the row's 27 outputs advance in parallel
// instead of one after the other (the warp-merge numeric pass spent ~85 % of its issue slots walking ballots: 7.8 ms at 256^3).
constexpr int kOwnStage = 12;      // entries of a B row staged per lane: 8 warps x 32 rows x 12 x 12 B = 36 KB per block
template <int SLOTS>
__global__ void __launch_bounds__(kWmWarps * 32) k_spgemm_owner(Csr A, Csr B, int shift, const int *__restrict__ c_rowptr,
                                                               int *__restrict__ c_col, double *__restrict__ c_val)
{
    constexpr unsigned kEnd = 0x7fffffffu;
    __shared__ unsigned s_bcol[kWmWarps][32 * kOwnStage];
    __shared__ double s_bval[kWmWarps][32 * kOwnStage];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int nwarps = gridDim.x * kWmWarps;
    for (int row = blockIdx.x * kWmWarps + w; row < A.rows; row += nwarps) {
        const int ks = A.rowptr[row], m = A.rowptr[row + 1] - ks;
        // ---- phase 1: sorted distinct columns, lane o keeps the o-th
        unsigned mycol = kEnd;
        {
            int q[SLOTS], qe[SLOTS];
            unsigned hc[SLOTS];
#pragma unroll
            for (int t = 0; t < SLOTS; ++t) {
                const int k = lane * SLOTS + t;
                q[t] = 0; qe[t] = 0; hc[t] = kEnd;
                if (k < m) {
                    const int j = A.col[ks + k] - shift;
                    q[t] = B.rowptr[j]; qe[t] = B.rowptr[j + 1];
                    if (q[t] < qe[t]) hc[t] = (unsigned)B.col[q[t]];
                }
            }
            int o = 0;
            while (true) {
                unsigned mine = kEnd;
#pragma unroll
                for (int t = 0; t < SLOTS; ++t) mine = min(mine, hc[t]);
                const unsigned cur = __reduce_min_sync(0xffffffffu, mine);
                if (cur == kEnd) break;
                if (lane == o) mycol = cur;
                ++o;
#pragma unroll
                for (int t = 0; t < SLOTS; ++t)
                    if (hc[t] == cur) { ++q[t]; hc[t] = q[t] < qe[t] ? (unsigned)B.col[q[t]] : kEnd; }
            }
        }
        // ---- phase 2: ordered accumulation by the owner lanes.  32 entries of the A row at a time: lane t first stages B row t
        // (up to kOwnStage entries) in the warp's shared-memory slab — 32 rows fetched in parallel, instead of one L2 round trip
        // per entry of the A row — then the warp walks the staged rows in order with broadcast reads.
        double sum = 0.0;
        bool have = false;
        unsigned *scol = s_bcol[w];
        double *sval = s_bval[w];
        for (int k0 = 0; k0 < m; k0 += 32) {
            const int kk = k0 + lane;
            int bq = 0, bqe = 0;
            double av = 0.0;
            if (kk < m) {
                const int j = A.col[ks + kk] - shift;
                bq = B.rowptr[j]; bqe = B.rowptr[j + 1];
                av = A.val[ks + kk];
            }
            __syncwarp();                        // the previous chunk's readers are done with the slab
            {
                const int len = min(bqe - bq, kOwnStage);
                for (int e = 0; e < len; ++e) { scol[lane * kOwnStage + e] = (unsigned)B.col[bq + e]; sval[lane * kOwnStage + e] = B.val[bq + e]; }
            }
            __syncwarp();
            const int cnt = min(32, m - k0);
            for (int t = 0; t < cnt; ++t) {
                const int e0 = __shfl_sync(0xffffffffu, bq, t), e1 = __shfl_sync(0xffffffffu, bqe, t);
                const double a = __shfl_sync(0xffffffffu, av, t);
                const int staged = min(e1 - e0, kOwnStage);
                for (int e = 0; e < staged; ++e) {
                    if (scol[t * kOwnStage + e] == mycol) {
                        const double pr = __dmul_rn(a, sval[t * kOwnStage + e]);
                        sum = have ? __dadd_rn(sum, pr) : pr;
                        have = true;
                    }
                }
                for (int e = e0 + staged; e < e1; ++e) {       // rows longer than the stage: the rest from global memory
                    if ((unsigned)B.col[e] == mycol) {
                        const double pr = __dmul_rn(a, B.val[e]);
                        sum = have ? __dadd_rn(sum, pr) : pr;
                        have = true;
                    }
                }
            }
        }
        if (mycol != kEnd) {
            const int o = c_rowptr[row] + lane;
            c_col[o] = (int)mycol;
            c_val[o] = sum;
        }
    }
}

template <int SLOTS>
static int run_wmerge(mgcg_ctx *c, const Csr &A, const Csr &B, int shift, DevCsr *C)
{
    cudaStream_t st = c->stream;
    int *cnt = nullptr;
    MGCG_CUDA(pool_alloc((void **)&cnt, sizeof(int) * ((size_t)A.rows + 1), st));
    MGCG_CUDA(cudaMemsetAsync(cnt + A.rows, 0, sizeof(int), st));
    int grid = (A.rows + kWmWarps - 1) / kWmWarps;
    if (grid > c->sm_count * 32) grid = c->sm_count * 32;
    k_spgemm_wmerge<false, SLOTS><<<grid, kWmWarps * 32, 0, st>>>(A, B, shift, cnt, nullptr, nullptr, nullptr);
    MGCG_LAUNCH_CHECK();
    C->rows = A.rows; C->cols = B.cols;
    MGCG_CUDA(pool_alloc((void **)&C->rowptr, sizeof(int) * ((size_t)A.rows + 1), st));
    int nnz = 0, maxcnt = 0;
    {
        // longest output row: at most 32 entries -> the owner-lane numeric pass
        size_t tb = 0;
        int *d_maxc = nullptr;
        void *tmp = nullptr;
        MGCG_CUDA(pool_alloc((void **)&d_maxc, sizeof(int), st));
        MGCG_CUDA(cub::DeviceReduce::Max(nullptr, tb, cnt, d_maxc, A.rows, st));
        MGCG_CUDA(pool_alloc(&tmp, tb > 0 ? tb : 1, st));
        MGCG_CUDA(cub::DeviceReduce::Max(tmp, tb, cnt, d_maxc, A.rows, st));
        MGCG_CUDA(cudaMemcpyAsync(&maxcnt, d_maxc, sizeof(int), cudaMemcpyDeviceToHost, st));
        const int s_ = scan_counts_checked(c, cnt, A.rows, C->rowptr, &nnz, "spgemm");     // synchronises
        pool_free(tmp, st); pool_free(d_maxc, st);
        pool_free(cnt, st);
        if (s_ != MGCG_OK) { pool_free(C->rowptr, st); C->rowptr = nullptr; return s_; }
    }
    C->nnz = nnz;
    MGCG_CUDA(pool_alloc((void **)&C->val, sizeof(double) * (size_t)(nnz > 0 ? nnz : 1), st));
    MGCG_CUDA(pool_alloc((void **)&C->col, sizeof(int) * (size_t)(nnz > 0 ? nnz : 1), st));
    // opt-in (MGCG_SPGEMM_OWNER=1): bit-identical, but measured level with the merge pass (256^3 fine-level R*(AP): 9.7 against 9.4 ms)
    const bool owner = maxcnt <= 32 && getenv("MGCG_SPGEMM_OWNER") != nullptr;
    if (getenv("MGCG_TRACE")) fprintf(stderr, "[mgcg spgemm] rows %d warp-merge kernel (%d lists per lane, %s numeric pass, longest output row %d) nnz %d grid %d\n",
                                      A.rows, SLOTS, owner ? "owner-lane" : "merge", maxcnt, nnz, grid);
    if (owner) k_spgemm_owner<SLOTS><<<grid, kWmWarps * 32, 0, st>>>(A, B, shift, C->rowptr, C->col, C->val);
    else k_spgemm_wmerge<true, SLOTS><<<grid, kWmWarps * 32, 0, st>>>(A, B, shift, nullptr, C->rowptr, C->col, C->val);
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

static int pow2_at_least(long long v) { int p = 32; while (p < v) p <<= 1; return p; }

template <int G>
static int run_hash(mgcg_ctx *c, const Csr &A, const Csr &B, int shift, int maxprod, DevCsr *C, bool *done)
{
    cudaStream_t st = c->stream;
    *done = false;
    // symbolic table (keys only): sized from the product-count bound when that is small, else grown x4 on overflow —
    // rows of R*(A*P) have thousands of products but only a few hundred distinct columns
    long long bound = maxprod < B.cols ? maxprod : B.cols;
    int cap_s = pow2_at_least(bound + bound / 2 + 1);
    if (cap_s > 512 && B.cols > 4096) cap_s = 512;   // (a small column space bounds the distinct columns tightly: keep it)
    int *cnt = nullptr;
    MGCG_CUDA(pool_alloc((void **)&cnt, sizeof(int) * ((size_t)A.rows + 3), st));
    const int cap_grid = c->sm_count * 16;
    int threads, grid;
    for (;; cap_s *= 4) {
        threads = 256;
        while (threads > G && threads > 32 && (size_t)(threads / G) * (cap_s + 1) * sizeof(int) > 96 * 1024) threads >>= 1;
        const size_t sm_s = (size_t)(threads / G) * (cap_s + 1) * sizeof(int);
        if (sm_s > 200 * 1024) { pool_free(cnt, st); return MGCG_OK; }   // too wide: caller falls back to spgemm_esc
        MGCG_CUDA(cudaFuncSetAttribute(k_spgemm_symbolic<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_s));
        MGCG_CUDA(cudaMemsetAsync(cnt + A.rows, 0, 3 * sizeof(int), st));
        const int gpb = threads / G;
        grid = (A.rows + gpb - 1) / gpb;
        if (grid > cap_grid) grid = cap_grid;
        k_spgemm_symbolic<G><<<grid, threads, sm_s, st>>>(A, B, shift, cap_s, cnt, cnt + A.rows + 2);
        MGCG_LAUNCH_CHECK();
        int ov = 0;
        MGCG_CUDA(cudaMemcpyAsync(&ov, cnt + A.rows + 2, sizeof(int), cudaMemcpyDeviceToHost, st));
        MGCG_CUDA(cudaStreamSynchronize(st));
        if (!ov) break;
    }
    // scan + max
    C->rows = A.rows; C->cols = B.cols;
    MGCG_CUDA(pool_alloc((void **)&C->rowptr, sizeof(int) * ((size_t)A.rows + 1), st));
    size_t tb = 0;
    int *d_max = cnt + A.rows + 1;
    MGCG_CUDA(cub::DeviceReduce::Max(nullptr, tb, cnt, d_max, A.rows, st));
    void *tmp = nullptr;
    MGCG_CUDA(pool_alloc(&tmp, tb, st));
    MGCG_CUDA(cub::DeviceReduce::Max(tmp, tb, cnt, d_max, A.rows, st));
    int h[2] = {0, 0};   // total nnz, max row length
    MGCG_CUDA(cudaMemcpyAsync(&h[1], d_max, sizeof(int), cudaMemcpyDeviceToHost, st));
    {
        const int s_ = scan_counts_checked(c, cnt, A.rows, C->rowptr, &h[0], "spgemm");   // synchronises
        pool_free(tmp, st); pool_free(cnt, st);
        if (s_ != MGCG_OK) { pool_free(C->rowptr, st); C->rowptr = nullptr; return s_; }
    }
    C->nnz = h[0];
    MGCG_CUDA(pool_alloc((void **)&C->val, sizeof(double) * (size_t)h[0], st));
    MGCG_CUDA(pool_alloc((void **)&C->col, sizeof(int) * (size_t)h[0], st));
    // numeric table: 1.5x the longest row
    const int cap_n = pow2_at_least((long long)h[1] + h[1] / 2 + 1);
    threads = 256;
    while (threads > G && (size_t)(threads / G) * (cap_n + 1) * 12 > 160 * 1024) threads >>= 1;
    if (threads < 32) threads = 32;
    const size_t sm_n = (size_t)(threads / G) * (cap_n + 1) * 12;
    if (sm_n > 200 * 1024) { C->free_all(); return MGCG_OK; }
    MGCG_CUDA(cudaFuncSetAttribute(k_spgemm_numeric<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_n));
    const int gpb_n = threads / G;
    grid = (A.rows + gpb_n - 1) / gpb_n;
    if (grid > cap_grid) grid = cap_grid;
    if (getenv("MGCG_TRACE"))
        fprintf(stderr, "[mgcg spgemm] rows %d G %d maxprod %d cap_sym %d cap_num %d (max row %d) nnz %d threads %d grid %d\n", A.rows, G, maxprod, cap_s, cap_n, h[1], h[0], threads, grid);
    k_spgemm_numeric<G><<<grid, threads, sm_n, st>>>(A, B, shift, cap_n, C->rowptr, C->col, C->val);
    MGCG_LAUNCH_CHECK();
    *done = true;
    return MGCG_OK;
}

// C = A * B.  Uses the shared-memory hash kernels; falls back to spgemm_esc for rows too wide for the tables.
int spgemm(mgcg_ctx *c, const Csr &A, const Csr &B, DevCsr *C, int col_shift)
{
    cudaStream_t st = c->stream;
    if ((A.rows <= 0 || A.nnz <= 0 || B.nnz <= 0) && col_shift == 0) return spgemm_esc(c, A, B, C);
    int *d_max = nullptr, hmax[2] = {0, 0};
    MGCG_CUDA(pool_alloc((void **)&d_max, 2 * sizeof(int), st));
    MGCG_CUDA(cudaMemsetAsync(d_max, 0, 2 * sizeof(int), st));
    int g = (A.rows + 255) / 256;
    if (g > c->sm_count * 16) g = c->sm_count * 16;
    k_spgemm_maxprod<<<g, 256, 0, st>>>(A, B, col_shift, d_max);
    k_spgemm_maxlen<<<g, 256, 0, st>>>(A, d_max + 1);
    MGCG_LAUNCH_CHECK();
    MGCG_CUDA(cudaMemcpyAsync(hmax, d_max, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    pool_free(d_max, st);
    const int maxprod = hmax[0], maxlen = hmax[1];
    if (!getenv("MGCG_SPGEMM_NO_MERGE")) {   // short A rows: per-thread k-way merge
        if (maxlen <= 8) return run_merge<8>(c, A, B, col_shift, C);
        if (maxlen <= 16) return run_merge<16>(c, A, B, col_shift, C);
        if (maxlen <= 32) return run_merge<32>(c, A, B, col_shift, C);
        const double avg_b0 = (double)B.nnz / (B.rows > 0 ? B.rows : 1);
        if (maxlen <= 32 * kWmSlots && avg_b0 <= 16.0) return run_wmerge<kWmSlots>(c, A, B, col_shift, C);
        if (maxlen <= 32 * kWmSlotsLong && avg_b0 <= 16.0) return run_wmerge<kWmSlotsLong>(c, A, B, col_shift, C);
    }
    const double avg_b = (double)B.nnz / (B.rows > 0 ? B.rows : 1);
    bool done = false;
    MGCG_CHECK(run_dense_rows(c, A, B, col_shift, maxprod, C, &done));
    if (done) return MGCG_OK;
    int s;
    if (avg_b <= 4.5) s = run_hash<4>(c, A, B, col_shift, maxprod, C, &done);
    else if (avg_b <= 10.0) s = run_hash<8>(c, A, B, col_shift, maxprod, C, &done);
    else if (avg_b <= 20.0) s = run_hash<16>(c, A, B, col_shift, maxprod, C, &done);
    else s = run_hash<32>(c, A, B, col_shift, maxprod, C, &done);
    MGCG_CHECK(s);
    if (done) return MGCG_OK;
    if (col_shift != 0) { set_error("spgemm: rows too wide for the hash tables on a shifted (sharded) product"); return MGCG_ERR_UNSUPPORTED; }
    return spgemm_esc(c, A, B, C);
}

}  // namespace mgcg

// ---- test hook (tests/test_gpu_mg.py): run the 64-bit-checked count scan on `rows` rows of `per_row` entries each
// without building a product of that size.  Returns the status; *nnz_out = total when it fits.
namespace mgcg {
__global__ void k_fill_int(int n, int v, int *a)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= n; i += gridDim.x * blockDim.x) a[i] = i < n ? v : 0;
}
}
extern "C" int mgcg_test_scan_counts(mgcg_ctx *c, int rows, int per_row, long long *nnz_out)
{
    using namespace mgcg;
    if (!c || rows <= 0 || per_row < 0) { set_error("mgcg_test_scan_counts: bad argument"); return MGCG_ERR_INVALID; }
    MGCG_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    int *cnt = nullptr, *rp = nullptr, nnz = 0;
    MGCG_CUDA(pool_alloc((void **)&cnt, sizeof(int) * ((size_t)rows + 1), st));
    MGCG_CUDA(pool_alloc((void **)&rp, sizeof(int) * ((size_t)rows + 1), st));
    k_fill_int<<<(rows + 256) / 256, 256, 0, st>>>(rows, per_row, cnt);
    const int s = scan_counts_checked(c, cnt, rows, rp, &nnz, "mgcg_test_scan_counts");
    pool_free(cnt, st); pool_free(rp, st);
    if (nnz_out) *nnz_out = (s == MGCG_OK) ? nnz : -1;
    return s;
}
