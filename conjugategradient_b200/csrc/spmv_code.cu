// spmv_code.cu — coded column indices: an internal layout behind the unchanged CSR interface (north_star subsystem 1: "an
// optional ... internal layout behind the unchanged CSR API"; the reference hands cusparseDcsrmv plain int32 column indices,
// Mgcg.cu:10-19).  On by default (mgcg_options.spmv_coding = MGCG_CODING_AUTO; MGCG_CODING_OFF keeps int32 columns).
//
// On a stencil matrix the DISTANCE col - row of an entry takes only a handful of values (7 for the 7-point operator, 27 for the
// 27-point one, whatever the grid size; a row-block shard adds a constant).  When a matrix has at most 256 distinct distances,
// its int32 column array (4 B per entry) is restated as one byte per entry — the rank of the entry's distance in the SORTED list
// of distinct distances — plus that list (<= 1 KB).  The SpMV kernel that understands the code (kernels_spmv.cu, k_spmv_tma with
// kCoded) stages 9 instead of 12 bytes per entry and rebuilds col = row + dict[code]: the same entries in the same storage order
// with the same x operands, so every row sum keeps its bits.  The CSR arrays the caller gave are not touched and stay the
// operands of everything else (hierarchy build, the uncoded kernels).
//
// Build: one pass collects the distinct distances (block-local hash set in shared memory, merged into a global one), one block
// sorts them, one pass writes the codes.  A matrix with more than 256 distinct distances (particle systems, aggregation
// operators) is simply left uncoded.
#include "common.cuh"
#include <limits.h>
#include <stdlib.h>

namespace mgcg {

constexpr int kCodeSlots = 1024;        // open-addressing slots of the distance sets (power of two, 4x the 256 codes)
constexpr int kCodeMax = 256;
constexpr int kCodeEmpty = INT_MIN;
constexpr int kCodeThreads = 256;

struct CodeState {
    int keys[kCodeSlots];
    int count, overflow, nseg, pad0;      // (read back by the host in one copy)
    int dict[kCodeMax];
    int ndict, pad;
    int seg_min[kSegMax];
    int seg_off[kCodeMax];
};

__device__ __forceinline__ unsigned code_hash(int d) { return ((unsigned)d * 2654435761u) >> 22; }   // 10 bits

// insert d into an open-addressing set; *count counts distinct keys; *overflow is raised beyond kCodeMax keys
__device__ __forceinline__ void code_insert(int *keys, int *count, int *overflow, int d)
{
    unsigned h = code_hash(d);
    for (int probe = 0; probe < kCodeSlots; ++probe) {
        const int cur = *(volatile int *)&keys[h];
        if (cur == d) return;
        if (cur == kCodeEmpty) {
            const int old = atomicCAS(&keys[h], kCodeEmpty, d);
            if (old == kCodeEmpty) {
                if (atomicAdd(count, 1) >= kCodeMax) *(volatile int *)overflow = 1;
                return;
            }
            if (old == d) return;
        }
        h = (h + 1) & (kCodeSlots - 1);
    }
    *(volatile int *)overflow = 1;
}

__global__ void __launch_bounds__(kCodeThreads) k_code_clear(CodeState *st)
{
    for (int i = threadIdx.x; i < kCodeSlots; i += kCodeThreads) st->keys[i] = kCodeEmpty;
    if (threadIdx.x < kCodeMax) st->dict[threadIdx.x] = 0;
    if (threadIdx.x == 0) { st->count = 0; st->overflow = 0; st->ndict = 0; st->nseg = 0; }
}

// Walk the entries of a tile of kCodeThreads consecutive rows with coalesced loads: entry k of the tile's contiguous slice belongs to
// the last row whose pointer is <= k (bisection in the tile's row pointers held in shared memory).  One thread per ROW reading its own
// 7..27 columns measured 5 ms per pass on the 27-point operator at 256^3 (every sector fetched once per 4-byte word it holds).
template <typename F>
__device__ __forceinline__ void walk_tile(const Csr &A, int tile, int *s_rp, F f)
{
    const int t = threadIdx.x;
    const int r0 = tile * kCodeThreads;
    const int nr = min(kCodeThreads, A.rows - r0);
    __syncthreads();
    if (t < nr) s_rp[t] = A.rowptr[r0 + t];
    if (t == 0) s_rp[nr] = A.rowptr[r0 + nr];
    __syncthreads();
    const int k0 = s_rp[0], k1 = s_rp[nr];
    for (int k = k0 + t; k < k1; k += kCodeThreads) {
        int lo = 0, hi = nr - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (s_rp[mid] <= k) lo = mid; else hi = mid - 1;
        }
        f(k, r0 + lo);
    }
}

__global__ void __launch_bounds__(kCodeThreads) k_code_collect(Csr A, CodeState *st)
{
    __shared__ int s_keys[kCodeSlots];
    __shared__ int s_rp[kCodeThreads + 1];
    __shared__ int s_count, s_over;
    const int t = threadIdx.x;
    for (int i = t; i < kCodeSlots; i += kCodeThreads) s_keys[i] = kCodeEmpty;
    if (t == 0) { s_count = 0; s_over = *(volatile int *)&st->overflow; }
    __syncthreads();
    const int *__restrict__ col = A.col;
    const int tiles = (A.rows + kCodeThreads - 1) / kCodeThreads;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        if (__syncthreads_or(*(volatile int *)&s_over)) break;        // block-uniform: walk_tile has barriers
        walk_tile(A, tile, s_rp, [&](int k, int row) { code_insert(s_keys, &s_count, &s_over, __ldcs(col + k) - row); });
    }
    __syncthreads();
    if (s_over) { if (t == 0) st->overflow = 1; return; }
    for (int i = t; i < kCodeSlots; i += kCodeThreads)
        if (s_keys[i] != kCodeEmpty) code_insert(st->keys, &st->count, &st->overflow, s_keys[i]);
}

// one block of kCodeSlots threads: sort the set ascending (bitonic, empties last) -> dict
__global__ void __launch_bounds__(kCodeSlots) k_code_sort(CodeState *st)
{
    __shared__ int s[kCodeSlots];
    const int t = threadIdx.x;
    const int v = st->keys[t];
    s[t] = (v == kCodeEmpty) ? INT_MAX : v;
    __syncthreads();
    for (int k = 2; k <= kCodeSlots; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            const int o = t ^ j;
            if (o > t) {
                const int a = s[t], b = s[o];
                const bool up = (t & k) == 0;
                if ((a > b) == up) { s[t] = b; s[o] = a; }
            }
            __syncthreads();
        }
    const int n = st->count < kCodeMax ? st->count : kCodeMax;
    if (t < kCodeMax) st->dict[t] = t < n ? s[t] : 0;
    if (t == 0) st->ndict = n;
}

// Group the sorted distances into runs: a run takes distances while they stay within kRunSpan of its first one.  One thread.
__global__ void k_code_runs(CodeState *st)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int n = st->ndict;
    int g = -1, first = 0;
    bool ok = n > 0 && st->overflow == 0;
    for (int i = 0; i < n && ok; ++i) {
        const int d = st->dict[i];
        if (g < 0 || (long long)d - first > kRunSpan) {
            if (++g >= kSegMax) { ok = false; break; }
            first = d;
            st->seg_min[g] = d;
        }
        st->seg_off[i] = g * 1024 + (d - first) + (first & 1);      // the run starts at the even index (r0 + first) & ~1, r0 even
    }
    st->nseg = ok ? g + 1 : 0;
}

__global__ void __launch_bounds__(kCodeThreads) k_code_encode(Csr A, const CodeState *st, unsigned char *code)
{
    __shared__ int s_dict[kCodeMax];
    __shared__ int s_rp[kCodeThreads + 1];
    const int t = threadIdx.x;
    s_dict[t] = st->dict[t];
    const int n = st->ndict;
    const int *__restrict__ col = A.col;
    const int tiles = (A.rows + kCodeThreads - 1) / kCodeThreads;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x)
        walk_tile(A, tile, s_rp, [&](int k, int row) {
            const int d = __ldcs(col + k) - row;
            int lo = 0, hi = n - 1;            // d is in the list: plain bisection
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (s_dict[mid] < d) lo = mid + 1; else hi = mid;
            }
            code[k] = (unsigned char)lo;
        });
}

// Restate A's columns as codes unless the option forbids it, when a kernel that reads them fits the matrix shape and the matrix has
// few enough distinct distances.  On success A->code / A->dict (and the run tables) point into `store` (kept across set-ups, grown on
// demand) and plan->stencil_rows is set when the STENCIL kernel applies; otherwise they stay null and the plain CSR kernels run.
// One host synchronisation (the verdict of the collect pass).
int spmv_code_build(mgcg_ctx *c, CodeStore *store, Csr *A, SpmvPlan *plan)
{
    A->code = nullptr; A->dict = nullptr; A->ndict = 0; A->seg_min = nullptr; A->seg_off = nullptr; A->nseg = 0;
    plan->stencil_rows = 0;
    const int min_rows = 1 << 15;          // below this the SpMV is launch-bound and the build would not pay
    static const bool env_off = [] { const char *e = getenv("MGCG_SPMV_CODED"); return e && e[0] == '0'; }();
    static const bool no_stencil = getenv("MGCG_NO_STENCIL") != nullptr;
    const int tile_rows = no_stencil ? 0 : spmv_plan_stencil_rows(*plan, A->rows, A->nnz);
    if (c->opt.spmv_coding == MGCG_CODING_OFF || env_off || A->rows < min_rows || A->nnz <= 0) return MGCG_OK;
    if (!spmv_plan_reads_code(*plan) && tile_rows == 0) return MGCG_OK;
    if (!store->state) MGCG_CUDA(pool_alloc(&store->state, sizeof(CodeState), c->stream));
    const size_t need = (((size_t)A->nnz + 15) & ~(size_t)15) + 16;     // whole 16-byte pieces: the TMA slices may run past the end
    if (store->cap < need) {
        if (store->buf) { pool_free(store->buf, c->stream); store->buf = nullptr; store->cap = 0; }
        MGCG_CUDA(pool_alloc((void **)&store->buf, need, c->stream));
        store->cap = need;
    }
    CodeState *st = (CodeState *)store->state;
    int grid = div_up(A->rows, kCodeThreads);
    if (grid > c->sm_count * 8) grid = c->sm_count * 8;
    k_code_clear<<<1, kCodeThreads, 0, c->stream>>>(st);
    k_code_collect<<<grid, kCodeThreads, 0, c->stream>>>(*A, st);
    k_code_sort<<<1, kCodeSlots, 0, c->stream>>>(st);
    k_code_runs<<<1, 32, 0, c->stream>>>(st);
    MGCG_LAUNCH_CHECK();
    int verdict[3] = {0, 1, 0};
    MGCG_CUDA(cudaMemcpyAsync(verdict, &st->count, sizeof(verdict), cudaMemcpyDeviceToHost, c->stream));
    MGCG_CUDA(cudaStreamSynchronize(c->stream));
    if (verdict[1] != 0 || verdict[0] > kCodeMax || verdict[0] <= 0) return MGCG_OK;     // too many distances: stay with CSR
    const bool stencil = tile_rows > 0 && verdict[2] > 0;
    if (!stencil && !spmv_plan_reads_code(*plan)) return MGCG_OK;                        // (a ROWS-shaped matrix with too many runs)
    MGCG_CUDA(cudaMemsetAsync(store->buf + (need - 32), 0, 32, c->stream));              // the slack past the last entry
    k_code_encode<<<grid, kCodeThreads, 0, c->stream>>>(*A, st, store->buf);
    MGCG_LAUNCH_CHECK();
    A->code = store->buf;
    A->dict = st->dict;
    A->ndict = verdict[0];
    if (stencil) { A->seg_min = st->seg_min; A->seg_off = st->seg_off; A->nseg = verdict[2]; plan->stencil_rows = tile_rows; }
    return MGCG_OK;
}

void spmv_code_free(CodeStore *store)
{
    if (store->buf) pool_release(store->buf);
    if (store->state) pool_release(store->state);
    store->buf = nullptr; store->state = nullptr; store->cap = 0;
}

}  // namespace mgcg
