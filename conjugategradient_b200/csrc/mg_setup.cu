// mg_setup.cu — multigrid hierarchy built ON THE DEVICE from the caller's CSR matrix.
// The reference has no multigrid at all ("Mgcg" is a name only, MgcgMain.cs:8); this is new work
// specified by BASELINE.json north_star item (3).  Parity is against the CPU twin oracle/mg_ref.c, which runs the
// same steps in the same order: every floating-point operation that defines the hierarchy (prolongator smoothing,
// Galerkin products, dense coarse inverse) is an explicit rounded multiply / add in a fixed order (no FMA
// contraction), so level operators can be compared bit-for-bit.
//
//   aggregates : geometric 2x2x2 blocks when a grid hint is given, else blocks of 8 consecutive rows
//                (== octants for Morton-sorted particles)
//   P          : plain aggregation T, or smoothed aggregation (I - w D^-1 A) T
//   A_coarse   : R (A P), R = P^T, both products by expand / stable radix sort / ordered segment sum (ESC)
//   coarsest   : dense inverse by in-place Gauss-Jordan (SPD, no pivoting)
// cub::DeviceRadixSort / DeviceScan are used for the sort and the scans of the SETUP only (not the solve path).
#include "mg.cuh"
#include <string.h>
#include <cub/cub.cuh>
#include <chrono>
#include <stdlib.h>

namespace mgcg {

typedef unsigned long long u64;

static thread_local cudaStream_t g_alloc_stream = nullptr;   // stream of the build in flight on this host thread
void set_alloc_stream(cudaStream_t st) { g_alloc_stream = st; }

template <class T>
static int dalloc(T **p, size_t n)
{
    *p = nullptr;
    if (n == 0) n = 1;
    MGCG_CUDA(pool_alloc((void **)p, n * sizeof(T), g_alloc_stream));
    return MGCG_OK;
}
static inline void tfree(void *p) { pool_free(p, g_alloc_stream); }

void DevCsr::free_all()
{
    pool_release(val);
    pool_release(rowptr);
    pool_release(col);
    val = nullptr; rowptr = nullptr; col = nullptr; rows = cols = nnz = 0;
}

static inline int grid_for(long long n, int threads = 256)
{
    long long g = (n + threads - 1) / threads;
    if (g < 1) g = 1;
    if (g > 1048576) g = 1048576;
    return (int)g;
}

// ---------------------------------------------------------------------------------------------------------
__global__ void k_extract_dinv(Csr A, int row_base, double *dinv)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.rows; i += gridDim.x * blockDim.x) {
        double d = 1.0;
        for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k)
            if (A.col[k] == i + row_base) { d = A.val[k]; break; }
        dinv[i] = 1.0 / d;
    }
}

__global__ void k_agg_geometric(int nx, int ny, int nz, int cx, int cy, int a, int *agg)
{
    const long long n = (long long)nx * ny * nz;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        int x = (int)(i % nx), y = (int)((i / nx) % ny), z = (int)(i / ((long long)nx * ny));
        agg[i] = (x / a) + cx * ((y / a) + cy * (z / a));
    }
}

__global__ void k_agg_blocks(int n, int block, int *agg)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) agg[i] = i / block;
}

// plain aggregation: P = T (one unit entry per row)
__global__ void k_build_T(int n, int row_base, const int *agg, double *val, int *rowptr, int *col)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= n; i += gridDim.x * blockDim.x) {
        rowptr[i] = i;
        if (i < n) { val[i] = 1.0; col[i] = agg[i + row_base]; }
    }
}

// smoothed aggregation, expansion of P = T - w D^-1 A T : per row i the entry (agg[i], 1) FIRST, then for every
// stored a_ik in storage order (agg[col_k], -(w*dinv_i)*a_ik).  Duplicates are summed later in this order.
__global__ void k_expand_sa(Csr A, int row_base, const int *agg, const double *dinv, double w, u64 *keys, double *vals)
{
    const int nz0 = A.rowptr[0];   // A may be a row sub-block of a larger CSR
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.rows; i += gridDim.x * blockDim.x) {
        size_t o = (size_t)(A.rowptr[i] - nz0) + i;
        keys[o] = ((u64)i << 32) | (unsigned)agg[i + row_base];
        vals[o] = 1.0;
        ++o;
        const double f = -__dmul_rn(w, dinv[i]);
        for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k, ++o) {
            keys[o] = ((u64)i << 32) | (unsigned)agg[A.col[k]];
            vals[o] = __dmul_rn(f, A.val[k]);
        }
    }
}

// The same expansion with one WARP per row, for the coarse levels (a few thousand rows of hundreds of entries: one thread
// per row left 6 blocks busy for a millisecond).  Every entry goes to the same position with the same value.
__global__ void k_expand_sa_warp(Csr A, int row_base, const int *agg, const double *dinv, double w, u64 *keys, double *vals)
{
    const int nz0 = A.rowptr[0];
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int i = warp; i < A.rows; i += nwarps) {
        const int ks = A.rowptr[i], ke = A.rowptr[i + 1];
        const size_t o0 = (size_t)(ks - nz0) + i;
        if (lane == 0) { keys[o0] = ((u64)i << 32) | (unsigned)agg[i + row_base]; vals[o0] = 1.0; }
        const double f = -__dmul_rn(w, dinv[i]);
        for (int k = ks + lane; k < ke; k += 32) {
            keys[o0 + 1 + (size_t)(k - ks)] = ((u64)i << 32) | (unsigned)agg[A.col[k]];
            vals[o0 + 1 + (size_t)(k - ks)] = __dmul_rn(f, A.val[k]);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// COO (key = row<<32 | col) -> CSR with ordered duplicate summation
__global__ void k_mark_heads(size_t count, const u64 *keys, int *head)
{
    for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < count; k += (size_t)gridDim.x * blockDim.x)
        head[k] = (k == 0 || keys[k] != keys[k - 1]) ? 1 : 0;
}
__global__ void k_write_heads(size_t count, const int *head, const int *idx, unsigned *pos, int uniq)
{
    for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < count; k += (size_t)gridDim.x * blockDim.x)
        if (head[k]) pos[idx[k]] = (unsigned)k;
    if (blockIdx.x == 0 && threadIdx.x == 0) pos[uniq] = (unsigned)count;
}
__global__ void k_compress(int uniq, const unsigned *pos, const u64 *keys, const double *vals,
                           double *oval, int *ocol, int *rowcnt)
{
    for (int o = blockIdx.x * blockDim.x + threadIdx.x; o < uniq; o += gridDim.x * blockDim.x) {
        unsigned k0 = pos[o], k1 = pos[o + 1];
        double s = vals[k0];
        for (unsigned k = k0 + 1; k < k1; ++k) s = __dadd_rn(s, vals[k]);
        oval[o] = s;
        u64 key = keys[k0];
        ocol[o] = (int)(key & 0xffffffffu);
        atomicAdd(&rowcnt[(int)(key >> 32)], 1);
    }
}

static int bits_for(unsigned v) { int b = 1; while ((v >> b) != 0 && b < 32) ++b; return b; }

// Sort `count` (key,val) pairs stably and compress to CSR rows [0,rows). keys/vals are consumed (scratch).
int coo_to_csr_sum(mgcg_ctx *c, int rows, int cols, size_t count, u64 *keys, double *vals, DevCsr *out)
{
    cudaStream_t st = c->stream;
    out->rows = rows; out->cols = cols;
    MGCG_CHECK(dalloc(&out->rowptr, (size_t)rows + 1));
    if (count == 0) {
        MGCG_CUDA(cudaMemsetAsync(out->rowptr, 0, sizeof(int) * ((size_t)rows + 1), st));
        out->nnz = 0;
        MGCG_CHECK(dalloc(&out->val, 1)); MGCG_CHECK(dalloc(&out->col, 1));
        return MGCG_OK;
    }
    if (count >= (1ull << 31)) { set_error("coo_to_csr_sum: %llu entries exceed the int32 index range", (unsigned long long)count); return MGCG_ERR_UNSUPPORTED; }
    u64 *keys2 = nullptr; double *vals2 = nullptr;
    MGCG_CHECK(dalloc(&keys2, count)); MGCG_CHECK(dalloc(&vals2, count));
    cub::DoubleBuffer<u64> dk(keys, keys2);
    cub::DoubleBuffer<double> dv(vals, vals2);
    const int end_bit = 32 + bits_for((unsigned)(rows > 0 ? rows - 1 : 0));
    size_t tmp_bytes = 0;
    MGCG_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, (long long)count, 0, end_bit, st));
    void *tmp = nullptr;
    MGCG_CUDA(pool_alloc(&tmp, tmp_bytes, st));
    MGCG_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, dk, dv, (long long)count, 0, end_bit, st));
    tfree(tmp);
    const u64 *sk = dk.Current(); const double *sv = dv.Current();

    int *head = nullptr, *idx = nullptr;
    MGCG_CHECK(dalloc(&head, count)); MGCG_CHECK(dalloc(&idx, count));
    k_mark_heads<<<grid_for((long long)count), 256, 0, st>>>(count, sk, head);
    MGCG_LAUNCH_CHECK();
    size_t sbytes = 0;
    MGCG_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, sbytes, head, idx, (long long)count, st));
    MGCG_CUDA(pool_alloc(&tmp, sbytes, st));
    MGCG_CUDA(cub::DeviceScan::ExclusiveSum(tmp, sbytes, head, idx, (long long)count, st));
    tfree(tmp);
    int last_idx = 0, last_head = 0;
    MGCG_CUDA(cudaMemcpyAsync(&last_idx, idx + count - 1, sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaMemcpyAsync(&last_head, head + count - 1, sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    const int uniq = last_idx + last_head;
    unsigned *pos = nullptr;
    MGCG_CHECK(dalloc(&pos, (size_t)uniq + 1));
    k_write_heads<<<grid_for((long long)count), 256, 0, st>>>(count, head, idx, pos, uniq);
    MGCG_LAUNCH_CHECK();
    out->nnz = uniq;
    MGCG_CHECK(dalloc(&out->val, (size_t)uniq)); MGCG_CHECK(dalloc(&out->col, (size_t)uniq));
    int *rowcnt = nullptr;
    MGCG_CHECK(dalloc(&rowcnt, (size_t)rows + 1));
    MGCG_CUDA(cudaMemsetAsync(rowcnt, 0, sizeof(int) * ((size_t)rows + 1), st));
    k_compress<<<grid_for(uniq), 256, 0, st>>>(uniq, pos, sk, sv, out->val, out->col, rowcnt);
    MGCG_LAUNCH_CHECK();
    MGCG_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, sbytes, rowcnt, out->rowptr, rows + 1, st));
    MGCG_CUDA(pool_alloc(&tmp, sbytes, st));
    MGCG_CUDA(cub::DeviceScan::ExclusiveSum(tmp, sbytes, rowcnt, out->rowptr, rows + 1, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    tfree(tmp);
    tfree(rowcnt); tfree(pos); tfree(head); tfree(idx); tfree(keys2); tfree(vals2);
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
// C = A * B by expand / sort / compress, in row chunks bounded by a product budget.
__global__ void k_spgemm_count(Csr A, Csr B, long long *cnt)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.rows; i += gridDim.x * blockDim.x) {
        long long s = 0;
        for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k) { int j = A.col[k]; s += B.rowptr[j + 1] - B.rowptr[j]; }
        cnt[i] = s;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) cnt[A.rows] = 0;
}
// largest `end` in (start, rows] with off[end]-off[start] <= budget (at least start+1)
__global__ void k_find_chunk_end(const long long *off, int rows, int start, long long budget, int *out_end, long long *out_cnt)
{
    int lo = start + 1, hi = rows;
    const long long base = off[start];
    while (lo < hi) { int mid = lo + (hi - lo + 1) / 2; if (off[mid] - base <= budget) lo = mid; else hi = mid - 1; }
    *out_end = lo; *out_cnt = off[lo] - base;
}
__global__ void k_spgemm_expand(Csr A, Csr B, const long long *off, int row0, int row1, u64 *keys, double *vals)
{
    for (int i = row0 + blockIdx.x * blockDim.x + threadIdx.x; i < row1; i += gridDim.x * blockDim.x) {
        size_t o = (size_t)(off[i] - off[row0]);
        for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k) {
            const int j = A.col[k];
            const double a = A.val[k];
            for (int q = B.rowptr[j]; q < B.rowptr[j + 1]; ++q, ++o) {
                keys[o] = ((u64)(unsigned)(i - row0) << 32) | (unsigned)B.col[q];
                vals[o] = __dmul_rn(a, B.val[q]);
            }
        }
    }
}
__global__ void k_add_offset(int n, int *a, int off)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) a[i] += off;
}

int spgemm_esc(mgcg_ctx *c, const Csr &A, const Csr &B, DevCsr *C)
{
    cudaStream_t st = c->stream;
    const long long budget = 1ll << 27;   // products per chunk (x 32 B of sort buffers)
    long long *cnt = nullptr, *off = nullptr;
    MGCG_CHECK(dalloc(&cnt, (size_t)A.rows + 1)); MGCG_CHECK(dalloc(&off, (size_t)A.rows + 1));
    k_spgemm_count<<<grid_for(A.rows), 256, 0, st>>>(A, B, cnt);
    MGCG_LAUNCH_CHECK();
    size_t sbytes = 0; void *tmp = nullptr;
    MGCG_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, sbytes, cnt, off, A.rows + 1, st));
    MGCG_CUDA(pool_alloc(&tmp, sbytes, st));
    MGCG_CUDA(cub::DeviceScan::ExclusiveSum(tmp, sbytes, cnt, off, A.rows + 1, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    tfree(tmp); tfree(cnt);
    int *d_end = nullptr; long long *d_cnt = nullptr;
    MGCG_CHECK(dalloc(&d_end, 1)); MGCG_CHECK(dalloc(&d_cnt, 1));

    std::vector<DevCsr> parts;
    std::vector<int> part_row0;
    int row0 = 0;
    while (row0 < A.rows) {
        int row1 = 0; long long count = 0;
        k_find_chunk_end<<<1, 1, 0, st>>>(off, A.rows, row0, budget, d_end, d_cnt);
        MGCG_LAUNCH_CHECK();
        MGCG_CUDA(cudaMemcpyAsync(&row1, d_end, sizeof(int), cudaMemcpyDeviceToHost, st));
        MGCG_CUDA(cudaMemcpyAsync(&count, d_cnt, sizeof(long long), cudaMemcpyDeviceToHost, st));
        MGCG_CUDA(cudaStreamSynchronize(st));
        u64 *keys = nullptr; double *vals = nullptr;
        MGCG_CHECK(dalloc(&keys, (size_t)count)); MGCG_CHECK(dalloc(&vals, (size_t)count));
        if (count > 0) {
            k_spgemm_expand<<<grid_for(row1 - row0, 128), 128, 0, st>>>(A, B, off, row0, row1, keys, vals);
            MGCG_LAUNCH_CHECK();
        }
        DevCsr part;
        MGCG_CHECK(coo_to_csr_sum(c, row1 - row0, B.cols, (size_t)count, keys, vals, &part));
        tfree(keys); tfree(vals);
        parts.push_back(part); part_row0.push_back(row0);
        row0 = row1;
    }
    tfree(off); tfree(d_end); tfree(d_cnt);
    // concatenate
    if (parts.size() == 1) { *C = parts[0]; C->rows = A.rows; C->cols = B.cols; return MGCG_OK; }
    long long total = 0;
    for (auto &p : parts) total += p.nnz;
    if (total >= (1ll << 31)) { set_error("spgemm_esc: coarse operator exceeds int32 non-zeros"); return MGCG_ERR_UNSUPPORTED; }
    C->rows = A.rows; C->cols = B.cols; C->nnz = (int)total;
    MGCG_CHECK(dalloc(&C->val, (size_t)total)); MGCG_CHECK(dalloc(&C->col, (size_t)total));
    MGCG_CHECK(dalloc(&C->rowptr, (size_t)A.rows + 1));
    int nz0 = 0;
    for (size_t i = 0; i < parts.size(); ++i) {
        DevCsr &p = parts[i];
        MGCG_CUDA(cudaMemcpyAsync(C->val + nz0, p.val, sizeof(double) * (size_t)p.nnz, cudaMemcpyDeviceToDevice, st));
        MGCG_CUDA(cudaMemcpyAsync(C->col + nz0, p.col, sizeof(int) * (size_t)p.nnz, cudaMemcpyDeviceToDevice, st));
        k_add_offset<<<grid_for(p.rows + 1), 256, 0, st>>>(p.rows + 1, p.rowptr, nz0);
        MGCG_LAUNCH_CHECK();
        MGCG_CUDA(cudaMemcpyAsync(C->rowptr + part_row0[i], p.rowptr, sizeof(int) * ((size_t)p.rows + 1), cudaMemcpyDeviceToDevice, st));
        nz0 += p.nnz;
    }
    MGCG_CUDA(cudaStreamSynchronize(st));
    for (auto &p : parts) p.free_all();
    return MGCG_OK;
}


// Transposition has no duplicates and its input is already ordered by row, so a STABLE sort by column alone (32-bit keys,
// only the bits the column count needs: 3 radix passes at 256^3 instead of 6 over 64-bit (col,row) keys) leaves every
// output row ordered by its column = the input row; the row pointers are read off the sorted keys.
struct RowVal { int row; int pad; double val; };

__global__ void k_transpose_expand32(Csr A, unsigned *keys, RowVal *rv)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.rows; i += gridDim.x * blockDim.x)
        for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k) {
            keys[k] = (unsigned)A.col[k];
            RowVal e; e.row = i; e.pad = 0; e.val = A.val[k];
            rv[k] = e;
        }
}
__global__ void k_transpose_finish(size_t count, int rows_t, const unsigned *__restrict__ keys, const RowVal *__restrict__ rv,
                                   int *__restrict__ rowptr, int *__restrict__ col, double *__restrict__ val)
{
    for (size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x; k < count; k += (size_t)gridDim.x * blockDim.x) {
        const int cur = (int)keys[k], prev = k > 0 ? (int)keys[k - 1] : -1;
        for (int r = prev + 1; r <= cur; ++r) rowptr[r] = (int)k;
        if (k == count - 1) for (int r = cur + 1; r <= rows_t; ++r) rowptr[r] = (int)count;
        col[k] = rv[k].row; val[k] = rv[k].val;
    }
}

int csr_transpose(mgcg_ctx *c, const Csr &A, DevCsr *T)
{
    cudaStream_t st = c->stream;
    const size_t count = (size_t)A.nnz;
    T->rows = A.cols; T->cols = A.rows; T->nnz = A.nnz;
    MGCG_CHECK(dalloc(&T->rowptr, (size_t)A.cols + 1));
    MGCG_CHECK(dalloc(&T->val, count)); MGCG_CHECK(dalloc(&T->col, count));
    if (count == 0) { MGCG_CUDA(cudaMemsetAsync(T->rowptr, 0, sizeof(int) * ((size_t)A.cols + 1), st)); return MGCG_OK; }
    unsigned *k1 = nullptr, *k2 = nullptr; RowVal *v1 = nullptr, *v2 = nullptr;
    MGCG_CHECK(dalloc(&k1, count)); MGCG_CHECK(dalloc(&k2, count)); MGCG_CHECK(dalloc(&v1, count)); MGCG_CHECK(dalloc(&v2, count));
    k_transpose_expand32<<<grid_for(A.rows), 256, 0, st>>>(A, k1, v1);
    MGCG_LAUNCH_CHECK();
    cub::DoubleBuffer<unsigned> dk(k1, k2);
    cub::DoubleBuffer<RowVal> dv(v1, v2);
    const int end_bit = bits_for((unsigned)(A.cols > 0 ? A.cols - 1 : 0));
    size_t tb = 0;
    MGCG_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, dk, dv, (long long)count, 0, end_bit, st));
    void *tmp = nullptr;
    MGCG_CUDA(pool_alloc(&tmp, tb, st));
    MGCG_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tb, dk, dv, (long long)count, 0, end_bit, st));   // LSD radix sort: stable
    k_transpose_finish<<<grid_for((long long)count), 256, 0, st>>>(count, A.cols, dk.Current(), dv.Current(), T->rowptr, T->col, T->val);
    MGCG_LAUNCH_CHECK();
    tfree(tmp); tfree(k1); tfree(k2); tfree(v1); tfree(v2);
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
// dense coarsest level: M = A^-1 by Gauss-Jordan without pivoting (SPD), one launch per pivot
__global__ void k_csr_to_dense(Csr A, double *M)
{
    const int n = A.rows;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k) M[(size_t)i * n + A.col[k]] = A.val[k];
}
// One launch per pivot, out of place (Min -> Mout, the buffers swap): every element recomputes its entry of the normalised
// pivot row instead of reading it from a second kernel's output, so nothing it reads is written in the same launch.
// Per element:  row k: m_kj / p (1 / p on the diagonal);  column k: -(m_ik * (1 / p));  else m_ij - m_ik * (m_kj / p).
__global__ void k_gj_step(int n, int k, const double *__restrict__ Min, double *__restrict__ Mout)
{
    const double p = Min[(size_t)k * n + k];
    const long long nn = (long long)n * n;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < nn; e += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(e / n), j = (int)(e % n);
        const double rkj = (j == k) ? 1.0 / p : Min[(size_t)k * n + j] / p;
        double v;
        if (i == k) v = rkj;
        else {
            const double ci = Min[(size_t)i * n + k];
            v = (j == k) ? -__dmul_rn(ci, rkj) : __dadd_rn(Min[e], -__dmul_rn(ci, rkj));
        }
        Mout[e] = v;
    }
}

// (Measured and dropped: the elimination in ONE block with the pivot row/column in shared memory and a block barrier per
// pivot.  Every update is an L2 round trip for the one SM doing the work: 2.0-3.0 ms at n = 216 against 1.85 ms with two
// launches per pivot.)
// BLOCKED Gauss-Jordan: kGjB pivots per pair of launches, the same arithmetic as k_gj_step element for element (so the inverse is
// bit-identical to the one-launch-per-pivot version and to the CPU twin, oracle/mg_ref.c dense_inverse).  For the pivots
// K = [k0, k0 + b) of a step, the K x K block evolves on its own; the b pivot rows and b pivot columns evolve pivot by pivot from
// the block's history (k_gjb_panels, which also records every normalised pivot row R[t][j] and pivot column C[i][t] as the
// one-pivot kernel would have seen them); every other element then takes its b updates in pivot order (k_gjb_update):
//     m_ij <- m_ij + (-(C[i][t] * R[t][j])),  t = 0 .. b-1.
// n = 1331: 84 x 2 launches instead of 1331 (measured with one launch per pivot: +17 ms of hierarchy build).
constexpr int kGjB = 16;
__global__ void __launch_bounds__(256) k_gjb_panels(int n, int k0, int b, const double *__restrict__ Min, double *__restrict__ Mout,
                                                    double *__restrict__ Rt, double *__restrict__ Ct)
{
    __shared__ double D[kGjB][kGjB + 1];        // the K x K block, evolving
    __shared__ double CK[kGjB][kGjB];           // CK[t][i]: column of pivot t inside the block, before the pivot (i != t)
    __shared__ double RK[kGjB][kGjB];           // RK[t][j]: normalised row of pivot t inside the block
    __shared__ double P[kGjB];
    const int tid = threadIdx.x, li = tid / kGjB, lj = tid % kGjB;
    if (li < b && lj < b) D[li][lj] = Min[(size_t)(k0 + li) * n + (k0 + lj)];
    __syncthreads();
    for (int t = 0; t < b; ++t) {
        double v = 0.0;
        if (li < b && lj < b) {
            const double p = D[t][t];
            const double rkj = (lj == t) ? 1.0 / p : D[t][lj] / p;
            if (li == t) v = rkj;
            else {
                const double ci = D[li][t];
                v = (lj == t) ? -__dmul_rn(ci, rkj) : __dadd_rn(D[li][lj], -__dmul_rn(ci, rkj));
            }
            if (li == t) { RK[t][lj] = rkj; if (lj == t) P[t] = p; }
            if (lj == t) CK[t][li] = D[li][t];
        }
        __syncthreads();
        if (li < b && lj < b) D[li][lj] = v;
        __syncthreads();
    }
    if (blockIdx.x == 0 && li < b && lj < b) Mout[(size_t)(k0 + li) * n + (k0 + lj)] = D[li][lj];
    // one thread per index g outside K: column g of the pivot rows, and row g of the pivot columns
    const int g = blockIdx.x * blockDim.x + tid;
    if (g >= n || (g >= k0 && g < k0 + b)) return;
    double u[kGjB], w[kGjB];
#pragma unroll
    for (int t = 0; t < kGjB; ++t) {
        u[t] = t < b ? Min[(size_t)(k0 + t) * n + g] : 0.0;      // M[K, g]
        w[t] = t < b ? Min[(size_t)g * n + (k0 + t)] : 0.0;      // M[g, K]
    }
#pragma unroll
    for (int t = 0; t < kGjB; ++t) {
        if (t < b) {
            const double r = u[t] / P[t];                         // normalised pivot row t at column g
            Rt[(size_t)t * n + g] = r;
            const double cg = w[t];                               // pivot column t at row g, before the pivot
            Ct[(size_t)g * kGjB + t] = cg;
#pragma unroll
            for (int i = 0; i < kGjB; ++i) {
                if (i < b && i != t) {
                    u[i] = __dadd_rn(u[i], -__dmul_rn(CK[t][i], r));
                    w[i] = __dadd_rn(w[i], -__dmul_rn(cg, RK[t][i]));
                }
            }
            u[t] = r;
            w[t] = -__dmul_rn(cg, RK[t][t]);
        }
    }
#pragma unroll
    for (int t = 0; t < kGjB; ++t) {
        if (t < b) { Mout[(size_t)(k0 + t) * n + g] = u[t]; Mout[(size_t)g * n + (k0 + t)] = w[t]; }
    }
}

__global__ void __launch_bounds__(256) k_gjb_update(int n, int k0, int b, const double *__restrict__ Min, double *__restrict__ Mout,
                                                    const double *__restrict__ Rt, const double *__restrict__ Ct)
{
    const int j = blockIdx.x * 64 + (threadIdx.x & 63);
    const int i0 = blockIdx.y * 16 + (threadIdx.x >> 6) * 4;
    if (j >= n || (j >= k0 && j < k0 + b)) return;
    double r[kGjB];
#pragma unroll
    for (int t = 0; t < kGjB; ++t) r[t] = t < b ? Rt[(size_t)t * n + j] : 0.0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int i = i0 + q;
        if (i >= n || (i >= k0 && i < k0 + b)) continue;
        double v = Min[(size_t)i * n + j];
#pragma unroll
        for (int t = 0; t < kGjB; ++t)
            if (t < b) v = __dadd_rn(v, -__dmul_rn(Ct[(size_t)i * kGjB + t], r[t]));
        Mout[(size_t)i * n + j] = v;
    }
}

static int dense_inverse(mgcg_ctx *c, const Csr &A, double **Minv)
{
    const int n = A.rows;
    cudaStream_t st = c->stream;
    double *M = nullptr, *M2 = nullptr;
    MGCG_CHECK(dalloc(&M, (size_t)n * n)); MGCG_CHECK(dalloc(&M2, (size_t)n * n));
    MGCG_CUDA(cudaMemsetAsync(M, 0, sizeof(double) * (size_t)n * n, st));
    k_csr_to_dense<<<grid_for(n), 256, 0, st>>>(A, M);
    MGCG_LAUNCH_CHECK();
    if (getenv("MGCG_GJ_UNBLOCKED")) {          // one launch per pivot (the round-1 path; tests compare the two bit for bit)
        for (int k = 0; k < n; ++k) {
            k_gj_step<<<grid_for((long long)n * n), 256, 0, st>>>(n, k, M, M2);
            double *t = M; M = M2; M2 = t;
        }
    } else {
        double *Rt = nullptr, *Ct = nullptr;
        MGCG_CHECK(dalloc(&Rt, (size_t)kGjB * n)); MGCG_CHECK(dalloc(&Ct, (size_t)kGjB * n));
        for (int k0 = 0; k0 < n; k0 += kGjB) {
            const int b = n - k0 < kGjB ? n - k0 : kGjB;
            k_gjb_panels<<<(n + 255) / 256, 256, 0, st>>>(n, k0, b, M, M2, Rt, Ct);
            k_gjb_update<<<dim3((unsigned)((n + 63) / 64), (unsigned)((n + 15) / 16)), 256, 0, st>>>(n, k0, b, M, M2, Rt, Ct);
            double *t = M; M = M2; M2 = t;
        }
        tfree(Rt); tfree(Ct);
    }
    MGCG_LAUNCH_CHECK();
    MGCG_CUDA(cudaStreamSynchronize(st));
    tfree(M2);
    *Minv = M;
    return MGCG_OK;
}


// ---------------------------------------------------------------------------------------------------------
// Smoothed prolongator WITHOUT the global sort, for matrices whose rows are short (row length + 1 <= MAXM): one thread per
// row keeps the row's (aggregate, value) pairs in registers, emits the distinct aggregates in ascending order and sums the
// duplicates of each in storage order — the same order and roundings as k_expand_sa + stable sort + k_compress (and as the
// CPU twin, oracle/mg_ref.c b_row), so P is bit-identical.  Replaces a 64-bit radix sort of nnz(A)+n pairs (12 ms of the
// 87 ms hierarchy build at 256^3) by two streaming passes over A.
template <int MAXM>
__device__ __forceinline__ int psa_load(const Csr &A, int i, int row_base, const int *__restrict__ agg, const double *__restrict__ dinv, double w,
                                        int (&cc)[MAXM], double (&vv)[MAXM], bool with_values)
{
    const int ks = A.rowptr[i], m = A.rowptr[i + 1] - ks + 1;
    const double f = with_values ? -__dmul_rn(w, dinv[i]) : 0.0;
#pragma unroll
    for (int k = 0; k < MAXM; ++k) {
        cc[k] = 0x7fffffff; vv[k] = 0.0;
        if (k == 0) { cc[0] = agg[i + row_base]; vv[0] = 1.0; }
        else if (k < m) {
            cc[k] = agg[A.col[ks + k - 1]];
            if (with_values) vv[k] = __dmul_rn(f, A.val[ks + k - 1]);
        }
    }
    return m;
}

template <int MAXM>
__global__ void __launch_bounds__(256) k_psa_count(Csr A, int row_base, const int *__restrict__ agg, int *__restrict__ cnt)
{
    int cc[MAXM]; double vv[MAXM];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.rows; i += gridDim.x * blockDim.x) {
        psa_load<MAXM>(A, i, row_base, agg, nullptr, 0.0, cc, vv, false);
        int distinct = 0;
#pragma unroll
        for (int k = 0; k < MAXM; ++k) {
            bool first = cc[k] != 0x7fffffff;
#pragma unroll
            for (int j = 0; j < k; ++j) first = first && (cc[j] != cc[k]);
            distinct += first ? 1 : 0;
        }
        cnt[i] = distinct;
    }
}

template <int MAXM>
__global__ void __launch_bounds__(256) k_psa_fill(Csr A, int row_base, const int *__restrict__ agg, const double *__restrict__ dinv, double w,
                                                  const int *__restrict__ p_rowptr, int *__restrict__ p_col, double *__restrict__ p_val)
{
    int cc[MAXM]; double vv[MAXM];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.rows; i += gridDim.x * blockDim.x) {
        psa_load<MAXM>(A, i, row_base, agg, dinv, w, cc, vv, true);
        const int o0 = p_rowptr[i], n_out = p_rowptr[i + 1] - o0;
        int last = -1;
        for (int o = 0; o < n_out; ++o) {
            int cur = 0x7fffffff;
#pragma unroll
            for (int k = 0; k < MAXM; ++k) if (cc[k] > last && cc[k] < cur) cur = cc[k];
            double sum = 0.0;
            bool have = false;
#pragma unroll
            for (int k = 0; k < MAXM; ++k)
                if (cc[k] == cur) { sum = have ? __dadd_rn(sum, vv[k]) : vv[k]; have = true; }
            p_col[o0 + o] = cur; p_val[o0 + o] = sum;
            last = cur;
        }
    }
}

__global__ void k_max_row(Csr A, int *out)
{
    int m = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.rows; i += gridDim.x * blockDim.x) m = max(m, A.rowptr[i + 1] - A.rowptr[i]);
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_down_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(out, m);
}

template <int MAXM>
static int psa_direct(mgcg_ctx *c, const Csr &A, int row_base, const int *agg, const double *dinv, double omega, int ncoarse, DevCsr *P)
{
    cudaStream_t st = c->stream;
    const int n = A.rows;
    int *cnt = nullptr;
    MGCG_CHECK(dalloc(&cnt, (size_t)n + 1));
    MGCG_CUDA(cudaMemsetAsync(cnt + n, 0, sizeof(int), st));
    int g = grid_for(n);
    if (g > c->sm_count * 32) g = c->sm_count * 32;
    k_psa_count<MAXM><<<g, 256, 0, st>>>(A, row_base, agg, cnt);
    MGCG_LAUNCH_CHECK();
    P->rows = n; P->cols = ncoarse;
    MGCG_CHECK(dalloc(&P->rowptr, (size_t)n + 1));
    int nnz = 0;
    {
        const int s_ = scan_counts_checked(c, cnt, n, P->rowptr, &nnz, "prolongator");
        tfree(cnt);
        if (s_ != MGCG_OK) return s_;
    }
    P->nnz = nnz;
    MGCG_CHECK(dalloc(&P->val, (size_t)nnz)); MGCG_CHECK(dalloc(&P->col, (size_t)nnz));
    k_psa_fill<MAXM><<<g, 256, 0, st>>>(A, row_base, agg, dinv, omega, P->rowptr, P->col, P->val);
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

// P = T (plain aggregation) or (I - w D^-1 A) T (smoothed) for the rows of A.  `agg` is indexed by GLOBAL row / column id
// (row i of A is global row i + row_base); columns of P are aggregate ids.
int build_prolongator(mgcg_ctx *c, const Csr &A, int row_base, const int *agg, const double *dinv, bool smooth, double omega,
                      int ncoarse, DevCsr *P)
{
    cudaStream_t st = c->stream;
    const int n = A.rows;
    if (!smooth) {
        P->rows = n; P->cols = ncoarse; P->nnz = n;
        MGCG_CHECK(dalloc(&P->val, (size_t)n)); MGCG_CHECK(dalloc(&P->col, (size_t)n)); MGCG_CHECK(dalloc(&P->rowptr, (size_t)n + 1));
        k_build_T<<<grid_for(n + 1), 256, 0, st>>>(n, row_base, agg, P->val, P->rowptr, P->col);
        MGCG_LAUNCH_CHECK();
        return MGCG_OK;
    }
    {   // short rows: no global sort
        int *d_max = nullptr, maxrow = 0;
        MGCG_CHECK(dalloc(&d_max, 1));
        MGCG_CUDA(cudaMemsetAsync(d_max, 0, sizeof(int), st));
        k_max_row<<<grid_for(n) > 4096 ? 4096 : grid_for(n), 256, 0, st>>>(A, d_max);
        MGCG_LAUNCH_CHECK();
        MGCG_CUDA(cudaMemcpyAsync(&maxrow, d_max, sizeof(int), cudaMemcpyDeviceToHost, st));
        MGCG_CUDA(cudaStreamSynchronize(st));
        tfree(d_max);
        if (maxrow + 1 <= 8) return psa_direct<8>(c, A, row_base, agg, dinv, omega, ncoarse, P);
        if (maxrow + 1 <= 16) return psa_direct<16>(c, A, row_base, agg, dinv, omega, ncoarse, P);
        if (maxrow + 1 <= 32) return psa_direct<32>(c, A, row_base, agg, dinv, omega, ncoarse, P);
    }
    const size_t count = (size_t)A.nnz + n;
    u64 *keys = nullptr; double *vals = nullptr;
    MGCG_CHECK(dalloc(&keys, count)); MGCG_CHECK(dalloc(&vals, count));
    if ((double)A.nnz >= 16.0 * n) k_expand_sa_warp<<<grid_for((long long)n * 32), 256, 0, st>>>(A, row_base, agg, dinv, omega, keys, vals);
    else k_expand_sa<<<grid_for(n), 256, 0, st>>>(A, row_base, agg, dinv, omega, keys, vals);
    MGCG_LAUNCH_CHECK();
    int s = coo_to_csr_sum(c, n, ncoarse, count, keys, vals, P);
    tfree(keys); tfree(vals);
    return s;
}

int extract_dinv(mgcg_ctx *c, const Csr &A, int row_base, double *dinv)
{
    k_extract_dinv<<<grid_for(A.rows), 256, 0, c->stream>>>(A, row_base, dinv);
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

// Gershgorin bound of the spectrum of D^-1 A: max over rows of sum_j |a_ij| |dinv_i|.  Row sums in storage order like the
// CPU twin (oracle/mg_ref.c mg_ref_set_chebyshev); the maximum is order-independent, so the result is bit-identical.
__global__ void k_gershgorin(Csr A, const double *__restrict__ dinv, unsigned long long *out)
{
    double m = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.rows; i += gridDim.x * blockDim.x) {
        double s = 0.0;
        for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k) s = __dadd_rn(s, fabs(A.val[k]));
        s = __dmul_rn(s, fabs(dinv[i]));
        m = fmax(m, s);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_down_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(out, (unsigned long long)__double_as_longlong(m));   // non-negative doubles order like their bits
}
int gershgorin_lmax(mgcg_ctx *c, const Csr &A, const double *dinv, double *out)
{
    unsigned long long *d = nullptr, h = 0;
    MGCG_CUDA(pool_alloc((void **)&d, sizeof(unsigned long long), c->stream));
    MGCG_CUDA(cudaMemsetAsync(d, 0, sizeof(unsigned long long), c->stream));
    k_gershgorin<<<grid_for(A.rows), 256, 0, c->stream>>>(A, dinv, d);
    MGCG_LAUNCH_CHECK();
    MGCG_CUDA(cudaMemcpyAsync(&h, d, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    MGCG_CUDA(cudaStreamSynchronize(c->stream));
    pool_release(d);
    memcpy(out, &h, sizeof(double));
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
static int alloc_level_vectors(MgLevel &L, bool level0)
{
    const size_t n = (size_t)L.A.rows;
    MGCG_CHECK(dalloc(&L.xa, n)); MGCG_CHECK(dalloc(&L.xb, n)); MGCG_CHECK(dalloc(&L.r, n));
    if (!level0) MGCG_CHECK(dalloc(&L.b, n));
    return MGCG_OK;
}

void mg_free(mgcg_ctx *c)
{
    if (!c->mg) return;
    Hierarchy *H = c->mg;
    for (size_t l = 0; l < H->L.size(); ++l) {
        MgLevel &L = H->L[l];
        L.Aown.free_all(); L.P.free_all(); L.R.free_all();
        spmv_code_free(&L.code);
        pool_release(L.agg);
        pool_release(L.dinv);
        pool_release(L.xa);
        pool_release(L.xb);
        pool_release(L.r);
        pool_release(L.d);
        if (l > 0) pool_release(L.b);
    }
    pool_release(H->coarse_inv);
    delete H;
    c->mg = nullptr;
}

// MGCG_TRACE=1: wall-clock of every phase of the hierarchy build on stderr (synchronises the stream at each mark)

int mg_build(mgcg_ctx *c)
{
    set_alloc_stream(c->stream);
    PhaseTrace tr(c->stream);
    mg_free(c);
    tr.mark(-1, "free-old");
    const mgcg_options &o = c->opt;
    if (o.mg_pre_sweeps < 1 || o.mg_post_sweeps < 0) { set_error("mg_build: mg_pre_sweeps must be >= 1, mg_post_sweeps >= 0"); return MGCG_ERR_INVALID; }
    if (o.mg_smoother != MGCG_SMOOTHER_JACOBI && o.mg_smoother != MGCG_SMOOTHER_RBGS && o.mg_smoother != MGCG_SMOOTHER_CHEBYSHEV) { set_error("mg_build: unknown smoother %d", o.mg_smoother); return MGCG_ERR_INVALID; }
    if (o.mg_smoother == MGCG_SMOOTHER_CHEBYSHEV && (o.mg_cheby_degree < 1 || o.mg_cheby_degree > 16 || !(o.mg_cheby_ratio > 1.0))) {
        set_error("mg_build: Chebyshev smoother needs 1 <= mg_cheby_degree <= 16 and mg_cheby_ratio > 1"); return MGCG_ERR_INVALID;
    }
    cudaStream_t st = c->stream;
    Hierarchy *H = new Hierarchy();
    c->mg = H;
    const int max_levels = o.mg_max_levels > 0 ? o.mg_max_levels : 12;
    const int coarse_size = o.mg_coarse_size > 0 ? o.mg_coarse_size : 1400;
    const bool hint = (o.nx > 0 && o.ny > 0 && o.nz > 0 && (long long)o.nx * o.ny * o.nz == c->A.rows);

    MgLevel L0;
    L0.A = c->A;
    L0.planA = c->planA;
    L0.dims[0] = hint ? o.nx : 0; L0.dims[1] = hint ? o.ny : 0; L0.dims[2] = hint ? o.nz : 0;
    H->L.push_back(L0);
    double nnz_sum = (double)c->A.nnz;

    for (int l = 0;; ++l) {
        MgLevel &L = H->L[l];
        const int n = L.A.rows;
        MGCG_CHECK(dalloc(&L.dinv, (size_t)n));
        k_extract_dinv<<<grid_for(n), 256, 0, st>>>(L.A, 0, L.dinv);
        MGCG_LAUNCH_CHECK();
        MGCG_CHECK(alloc_level_vectors(L, l == 0));
        if (n <= coarse_size || l + 1 >= max_levels) break;
        if (o.mg_smoother == MGCG_SMOOTHER_CHEBYSHEV) {
            L.cheb = o.mg_cheby_degree;
            MGCG_CHECK(gershgorin_lmax(c, L.A, L.dinv, &L.lmax));
            MGCG_CHECK(dalloc(&L.d, (size_t)n));
            if (tr.on) fprintf(stderr, "[mgcg setup] level %d smoother: Chebyshev degree %d on [%.4f, %.4f]\n", l, L.cheb, L.lmax / o.mg_cheby_ratio, L.lmax);
        }
        if (o.mg_smoother == MGCG_SMOOTHER_RBGS && L.dims[0] > 0 && L.planA.kernel == MGCG_SPMV_STREAM) {
            // red-black Gauss-Seidel where the level's operator is two-colourable by cell parity, weighted Jacobi elsewhere
            int two = 0;
            MGCG_CHECK(check_two_colourable(c, L.A, L.dims[0], L.dims[1], &two));
            L.rb = two != 0;
            if (tr.on) fprintf(stderr, "[mgcg setup] level %d smoother: %s\n", l, L.rb ? "red-black Gauss-Seidel" : "weighted Jacobi (not two-colourable)");
        }

        // ---- aggregates
        int nc;
        int cd[3] = {0, 0, 0};
        MGCG_CHECK(dalloc(&L.agg, (size_t)n));
        // aggregation: AUTO = geometric blocks with a grid hint, strength filter + MIS-2 without; BLOCKS = 8 consecutive rows
        // without a hint (== cells of cell-sorted particles; the round-1 rule); MIS2 = algebraic even with a hint
        const bool geometric = L.dims[0] > 0 && o.mg_aggregation != MGCG_AGG_MIS2;
        const bool filtered = (o.mg_aggregation == MGCG_AGG_MIS2) || (L.dims[0] <= 0 && o.mg_aggregation != MGCG_AGG_BLOCKS);
        DevCsr F;
        double *dinvF = nullptr;
        if (geometric) {
            int a = (l == 0 || o.mg_aggregate_size_coarse <= 0) ? o.mg_aggregate_size : o.mg_aggregate_size_coarse;
            if (a < 2) a = 2;
            cd[0] = (L.dims[0] + a - 1) / a; cd[1] = (L.dims[1] + a - 1) / a; cd[2] = (L.dims[2] + a - 1) / a;
            nc = cd[0] * cd[1] * cd[2];
            k_agg_geometric<<<grid_for(n), 256, 0, st>>>(L.dims[0], L.dims[1], L.dims[2], cd[0], cd[1], a, L.agg);
        } else if (!filtered) {
            nc = (n + 7) / 8;
            k_agg_blocks<<<grid_for(n), 256, 0, st>>>(n, 8, L.agg);
        } else {
            const double theta = o.mg_strength_theta > 0.0 ? o.mg_strength_theta : 0.5;
            int rounds = 0;
            MGCG_CHECK(strength_filter(c, L.A, theta, &F, &dinvF));
            tr.mark(l, "filter");
            int s_ = mis2_aggregate(c, F.view(), L.agg, &nc, &rounds);
            if (s_ != MGCG_OK) { F.free_all(); pool_release(dinvF); return s_; }
            if (tr.on) fprintf(stderr, "[mgcg setup] level %d MIS-2: %d rows -> %d aggregates in %d rounds, %.1f strong entries per row\n",
                               l, n, nc, rounds, (double)(F.nnz - n) / n);
        }
        MGCG_LAUNCH_CHECK();
        if (nc >= n) { F.free_all(); pool_release(dinvF); break; }   // cannot coarsen further
        tr.mark(l, "aggregate");

        // ---- prolongator (smoothed with the level operator, or with its strength-filtered version under MIS-2)
        const bool smooth_here = o.mg_smoothed && (o.mg_smoothed_levels <= 0 || l < o.mg_smoothed_levels);
        {
            const int s_ = filtered ? build_prolongator(c, F.view(), 0, L.agg, dinvF, smooth_here, o.mg_prolong_omega, nc, &L.P)
                                    : build_prolongator(c, L.A, 0, L.agg, L.dinv, smooth_here, o.mg_prolong_omega, nc, &L.P);
            F.free_all(); pool_release(dinvF);
            if (s_ != MGCG_OK) return s_;
        }
        tr.mark(l, "P");
        MGCG_CHECK(csr_transpose(c, L.P.view(), &L.R));
        tr.mark(l, "R=P^T");

        // ---- Galerkin coarse operator  A_c = R (A P)
        DevCsr AP, Ac;
        MGCG_CHECK(spgemm(c, L.A, L.P.view(), &AP));
        tr.mark(l, "A*P");
        int s = spgemm(c, L.R.view(), AP.view(), &Ac);
        tr.mark(l, "R*(AP)");
        AP.free_all();
        MGCG_CHECK(s);

        L.planP = spmv_make_plan(c, L.P.rows, L.P.nnz, MGCG_SPMV_AUTO);
        L.planR = spmv_make_plan(c, L.R.rows, L.R.nnz, MGCG_SPMV_AUTO);

        MgLevel N;
        N.Aown = Ac;
        N.A = Ac.view();
        N.planA = spmv_make_plan(c, Ac.rows, Ac.nnz, MGCG_SPMV_AUTO);
        MGCG_CHECK(spmv_code_build(c, &N.code, &N.A, &N.planA));     // a Galerkin operator on a grid is a stencil too (27-point on level 1)
        N.dims[0] = cd[0]; N.dims[1] = cd[1]; N.dims[2] = cd[2];
        nnz_sum += (double)Ac.nnz;
        H->L.push_back(N);   // invalidates L
    }
    // ---- coarsest level: dense inverse
    MgLevel &Lc = H->L.back();
    if (Lc.A.rows > 4096) { set_error("mg_build: coarsest level has %d rows (> 4096): raise mg_max_levels", Lc.A.rows); return MGCG_ERR_UNSUPPORTED; }
    H->coarse_n = Lc.A.rows;
    tr.mark((int)H->L.size() - 1, "pre-dense");
    MGCG_CHECK(dense_inverse(c, Lc.A, &H->coarse_inv));
    tr.mark((int)H->L.size() - 1, "dense-inverse");
    H->op_complexity = nnz_sum / (double)(c->A.nnz > 0 ? c->A.nnz : 1);
    mg_tail_plan(c);
    MGCG_CUDA(cudaStreamSynchronize(st));
    return MGCG_OK;
}

double *mg_level0_first_x(mgcg_ctx *c) { return c->mg ? c->mg->L[0].xa : nullptr; }
double mg_level0_first_omega(mgcg_ctx *c)
{
    if (c->mg && c->mg->L[0].cheb) { double c1, c2; cheby_coeffs(c->mg->L[0].lmax, c->opt.mg_cheby_ratio, 0, &c1, &c2); return c2; }
    return (c->mg && c->mg->L[0].rb) ? 1.0 : c->opt.mg_omega;
}
const double *mg_level0_dinv(mgcg_ctx *c) { return c->mg ? c->mg->L[0].dinv : nullptr; }

}  // namespace mgcg

// ---------------------------------------------------------------------------------------------------------
// introspection exports (tests: hierarchy parity against the CPU twin)
using namespace mgcg;

extern "C" int mgcg_mg_levels(mgcg_ctx *c) { return (c && c->mg) ? (int)c->mg->L.size() : 1; }

extern "C" int mgcg_mg_level_size(mgcg_ctx *c, int level, int *rows, int *nnz)
{
    if (!c || !c->mg || level < 0 || level >= (int)c->mg->L.size()) { set_error("mgcg_mg_level_size: no such level"); return MGCG_ERR_INVALID; }
    *rows = c->mg->L[level].A.rows; *nnz = c->mg->L[level].A.nnz;
    return MGCG_OK;
}

extern "C" int mgcg_mg_level_get(mgcg_ctx *c, int level, double *elements, int *row_offsets, int *column_indeces)
{
    if (!c || !c->mg || level < 0 || level >= (int)c->mg->L.size()) { set_error("mgcg_mg_level_get: no such level"); return MGCG_ERR_INVALID; }
    const Csr &A = c->mg->L[level].A;
    MGCG_CUDA(cudaSetDevice(c->device));
    MGCG_CUDA(cudaMemcpy(elements, A.val, sizeof(double) * (size_t)A.nnz, cudaMemcpyDeviceToHost));
    MGCG_CUDA(cudaMemcpy(column_indeces, A.col, sizeof(int) * (size_t)A.nnz, cudaMemcpyDeviceToHost));
    MGCG_CUDA(cudaMemcpy(row_offsets, A.rowptr, sizeof(int) * ((size_t)A.rows + 1), cudaMemcpyDeviceToHost));
    return MGCG_OK;
}

extern "C" int mgcg_mg_level_aggregates(mgcg_ctx *c, int level, int *agg)
{
    if (!c || !c->mg || level < 0 || level + 1 >= (int)c->mg->L.size()) { set_error("mgcg_mg_level_aggregates: no such level"); return MGCG_ERR_INVALID; }
    const MgLevel &L = c->mg->L[level];
    MGCG_CUDA(cudaSetDevice(c->device));
    MGCG_CUDA(cudaMemcpy(agg, L.agg, sizeof(int) * (size_t)L.A.rows, cudaMemcpyDeviceToHost));
    return MGCG_OK;
}
