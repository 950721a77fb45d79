// generate_mps.cu — particle (MPS) pressure-Poisson matrix assembled in device memory: neighbour search by cell list -> CSR.
// SURVEY.md §8(f)2 ("MPS neighbour search (cell list) -> CSR directly in device memory").  The reference receives this matrix
// from its particle-method caller (at most 160 entries per row, Mgcg/cuBlas/Mgcg/MgcgMain.cs:20); the semantics here are the
// ones written down in oracle/mps_ref.c and must match it bit for bit:
//   cells of edge re on the bounding box, id = cx + ncx*(cy + ncy*cz); particles numbered in cell order (stable);
//   neighbours r < re with r = sqrt((dx*dx + dy*dy) + dz*dz); a_ij = -(re/r - 1); a_ii = sum_j w_ij (ascending j) + shift;
//   rows: diagonal first (MgcgMain.cs:59-60), then ascending columns — which the (dz, dy) scan over contiguous x-runs of
//   cell-ordered particles produces without a per-row sort.
// STATUS: NOT YET RUN ON A GPU (written after the round's GPU budget was spent) — see tests/test_gpu_mps.py.
#include "common.cuh"
#include <cub/cub.cuh>
#include <math.h>

namespace mgcg {

struct MpsGrid { double x0, y0, z0, re; int ncx, ncy, ncz; };

__global__ void __launch_bounds__(256) k_mps_bbox(int n, const double *__restrict__ pos, double *__restrict__ part /* [grid][6] */)
{
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
#pragma unroll
        for (int d = 0; d < 3; ++d) { const double v = pos[3 * (size_t)i + d]; lo[d] = fmin(lo[d], v); hi[d] = fmax(hi[d], v); }
    __shared__ double s[6][8];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        for (int o = 16; o > 0; o >>= 1) {
            lo[d] = fmin(lo[d], __shfl_down_sync(0xffffffffu, lo[d], o));
            hi[d] = fmax(hi[d], __shfl_down_sync(0xffffffffu, hi[d], o));
        }
        if ((threadIdx.x & 31) == 0) { s[d][threadIdx.x >> 5] = lo[d]; s[3 + d][threadIdx.x >> 5] = hi[d]; }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        double v = s[threadIdx.x][0];
        for (int w = 1; w < 8; ++w) v = threadIdx.x < 3 ? fmin(v, s[threadIdx.x][w]) : fmax(v, s[threadIdx.x][w]);
        part[6 * (size_t)blockIdx.x + threadIdx.x] = v;
    }
}

__device__ __forceinline__ void mps_cell_xyz(const MpsGrid &g, const double *p, int &cx, int &cy, int &cz)
{
    cx = (int)floor(__ddiv_rn(__dsub_rn(p[0], g.x0), g.re));
    cy = (int)floor(__ddiv_rn(__dsub_rn(p[1], g.y0), g.re));
    cz = (int)floor(__ddiv_rn(__dsub_rn(p[2], g.z0), g.re));
}

__global__ void __launch_bounds__(256) k_mps_cell_keys(int n, MpsGrid g, const double *__restrict__ pos, unsigned *__restrict__ key,
                                                       int *__restrict__ idx)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        int cx, cy, cz;
        mps_cell_xyz(g, pos + 3 * (size_t)i, cx, cy, cz);
        key[i] = (unsigned)(cx + g.ncx * (cy + g.ncy * cz));
        idx[i] = i;
    }
}

__global__ void __launch_bounds__(256) k_mps_gather(int n, const int *__restrict__ order, const double *__restrict__ pos,
                                                    double *__restrict__ sorted)
{
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        const size_t s = 3 * (size_t)order[k];
        sorted[3 * (size_t)k] = pos[s]; sorted[3 * (size_t)k + 1] = pos[s + 1]; sorted[3 * (size_t)k + 2] = pos[s + 2];
    }
}

// cell_start[c] = first particle number of cell c (particles are in cell order); cell_start[ncells] = n.  *bad = 1 if not ordered.
__global__ void __launch_bounds__(256) k_mps_cell_start(int n, MpsGrid g, const double *__restrict__ pos, int ncells,
                                                        int *__restrict__ cell_start, int *bad)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= n; i += gridDim.x * blockDim.x) {
        int c0 = -1, c1 = ncells;
        int cx, cy, cz;
        if (i > 0) { mps_cell_xyz(g, pos + 3 * (size_t)(i - 1), cx, cy, cz); c0 = cx + g.ncx * (cy + g.ncy * cz); }
        if (i < n) { mps_cell_xyz(g, pos + 3 * (size_t)i, cx, cy, cz); c1 = cx + g.ncx * (cy + g.ncy * cz); }
        if (c1 < c0) { *bad = 1; continue; }
        for (int c = c0 + 1; c <= c1; ++c) cell_start[c] = i;      // the cells (c0, c1] begin at particle i (empty ones too)
    }
}

// One thread per particle.  FILL = false: counts (row length incl. the diagonal) into cnt[i]; FILL = true: writes the row.
template <bool FILL>
__global__ void __launch_bounds__(128) k_mps_rows(int n, MpsGrid g, const double *__restrict__ pos, const int *__restrict__ cell_start,
                                                  double shift, int max_per_row, int *__restrict__ cnt, const int *__restrict__ rowptr,
                                                  double *__restrict__ val, int *__restrict__ col, int *too_long)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n + (FILL ? 0 : 1); i += gridDim.x * blockDim.x) {
        if (!FILL && i == n) { cnt[n] = 0; continue; }
        const double px = pos[3 * (size_t)i], py = pos[3 * (size_t)i + 1], pz = pos[3 * (size_t)i + 2];
        const double p[3] = {px, py, pz};
        int cx, cy, cz;
        mps_cell_xyz(g, p, cx, cy, cz);
        const int xa = cx > 0 ? cx - 1 : 0, xb = cx < g.ncx - 1 ? cx + 1 : g.ncx - 1;
        long long o = FILL ? (long long)rowptr[i] + 1 : 1;
        double diag = 0.0;
        for (int dz = -1; dz <= 1; ++dz) {
            const int zz = cz + dz;
            if (zz < 0 || zz >= g.ncz) continue;
            for (int dy = -1; dy <= 1; ++dy) {
                const int yy = cy + dy;
                if (yy < 0 || yy >= g.ncy) continue;
                const long long base = (long long)g.ncx * (yy + (long long)g.ncy * zz);
                const int j0 = cell_start[base + xa], j1 = cell_start[base + xb + 1];
                for (int j = j0; j < j1; ++j) {
                    if (j == i) continue;
                    const double dx = __dsub_rn(px, pos[3 * (size_t)j]), dyv = __dsub_rn(py, pos[3 * (size_t)j + 1]),
                                 dzv = __dsub_rn(pz, pos[3 * (size_t)j + 2]);
                    const double r = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dyv, dyv)), __dmul_rn(dzv, dzv)));
                    if (!(r < g.re)) continue;
                    if (FILL) {
                        const double w = __dsub_rn(__ddiv_rn(g.re, r), 1.0);
                        diag = __dadd_rn(diag, w);
                        val[o] = -w; col[o] = j;
                    }
                    ++o;
                }
            }
        }
        if (FILL) { val[rowptr[i]] = __dadd_rn(diag, shift); col[rowptr[i]] = i; }
        else {
            cnt[i] = (int)o;
            if (o > max_per_row) *too_long = 1;
        }
    }
}

static int mps_grid_blocks(long long threads, int per, int sm)
{
    long long g = (threads + per - 1) / per;
    if (g > (long long)sm * 32) g = (long long)sm * 32;
    return (int)(g < 1 ? 1 : g);
}

// bounding box -> grid (one small D2H); < 0 on error
static int mps_make_grid(mgcg_ctx *c, int n, const double *d_pos, double re, MpsGrid *g)
{
    cudaStream_t st = c->stream;
    const int blocks = mps_grid_blocks(n, 256, c->sm_count) > 1024 ? 1024 : mps_grid_blocks(n, 256, c->sm_count);
    double *d_part = nullptr;
    MGCG_CUDA(pool_alloc((void **)&d_part, sizeof(double) * 6 * (size_t)blocks, st));
    k_mps_bbox<<<blocks, 256, 0, st>>>(n, d_pos, d_part);
    std::vector<double> h(6 * (size_t)blocks);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(h.data(), d_part, sizeof(double) * h.size(), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    pool_release(d_part);
    if (e != cudaSuccess) { set_error("mps bounding box: %s", cudaGetErrorString(e)); return MGCG_ERR_CUDA; }
    double lo[3] = {h[0], h[1], h[2]}, hi[3] = {h[3], h[4], h[5]};
    for (int b = 1; b < blocks; ++b)
        for (int d = 0; d < 3; ++d) {
            lo[d] = lo[d] < h[6 * (size_t)b + d] ? lo[d] : h[6 * (size_t)b + d];
            hi[d] = hi[d] > h[6 * (size_t)b + 3 + d] ? hi[d] : h[6 * (size_t)b + 3 + d];
        }
    g->x0 = lo[0]; g->y0 = lo[1]; g->z0 = lo[2]; g->re = re;
    g->ncx = (int)floor((hi[0] - lo[0]) / re) + 1;
    g->ncy = (int)floor((hi[1] - lo[1]) / re) + 1;
    g->ncz = (int)floor((hi[2] - lo[2]) / re) + 1;
    const long long nc = (long long)g->ncx * g->ncy * g->ncz;
    if (!(nc > 0) || nc >= (1ll << 31) - 1) { set_error("mps grid: %lld cells", nc); return MGCG_ERR_UNSUPPORTED; }
    return MGCG_OK;
}

// cell_start (pool block of ncells + 1 ints, caller releases) for positions already in cell order
static int mps_cell_start(mgcg_ctx *c, int n, const MpsGrid &g, const double *d_sorted, int **d_cell_start)
{
    cudaStream_t st = c->stream;
    const int ncells = g.ncx * g.ncy * g.ncz;
    int *cs = nullptr, *d_bad = nullptr;
    MGCG_CUDA(pool_alloc((void **)&cs, sizeof(int) * ((size_t)ncells + 1), st));
    MGCG_CUDA(pool_alloc((void **)&d_bad, sizeof(int), st));
    MGCG_CUDA(cudaMemsetAsync(d_bad, 0, sizeof(int), st));
    k_mps_cell_start<<<mps_grid_blocks((long long)n + 1, 256, c->sm_count), 256, 0, st>>>(n, g, d_sorted, ncells, cs, d_bad);
    int bad = 0;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    pool_release(d_bad);
    if (e != cudaSuccess) { pool_release(cs); set_error("mps cell list: %s", cudaGetErrorString(e)); return MGCG_ERR_CUDA; }
    if (bad) { pool_release(cs); set_error("mps: positions are not in cell order (call mgcg_mps_order first)"); return MGCG_ERR_INVALID; }
    *d_cell_start = cs;
    return MGCG_OK;
}

}  // namespace mgcg

using namespace mgcg;

// Number the particles in cell order: d_order[k] = input index of particle k, d_pos_sorted = positions in that order.
// Returns the number of cells (> 0) or a negative status.
extern "C" long long mgcg_mps_order(mgcg_ctx *c, int n, const double *d_pos, double re, int *d_order, double *d_pos_sorted)
{
    if (!c || n <= 0 || !(re > 0) || !d_pos || !d_order || !d_pos_sorted) { set_error("mgcg_mps_order: bad argument"); return -MGCG_ERR_INVALID; }
    if (cudaSetDevice(c->device) != cudaSuccess) { set_error("mgcg_mps_order: cudaSetDevice failed"); return -MGCG_ERR_CUDA; }
    cudaStream_t st = c->stream;
    MpsGrid g;
    int s = mps_make_grid(c, n, d_pos, re, &g);
    if (s != MGCG_OK) return -s;
    const int ncells = g.ncx * g.ncy * g.ncz;
    unsigned *k1 = nullptr, *k2 = nullptr; int *i1 = nullptr;
    void *tmp = nullptr;
    size_t tb = 0;
    int bits = 1; while ((ncells - 1) >> bits && bits < 32) ++bits;
    cudaError_t e = pool_alloc((void **)&k1, sizeof(unsigned) * (size_t)n, st);
    if (e == cudaSuccess) e = pool_alloc((void **)&k2, sizeof(unsigned) * (size_t)n, st);
    if (e == cudaSuccess) e = pool_alloc((void **)&i1, sizeof(int) * (size_t)n, st);
    if (e == cudaSuccess) {
        k_mps_cell_keys<<<mps_grid_blocks(n, 256, c->sm_count), 256, 0, st>>>(n, g, d_pos, k1, i1);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cub::DeviceRadixSort::SortPairs(nullptr, tb, k1, k2, i1, d_order, n, 0, bits, st);
    if (e == cudaSuccess) e = pool_alloc(&tmp, tb, st);
    if (e == cudaSuccess) e = cub::DeviceRadixSort::SortPairs(tmp, tb, k1, k2, i1, d_order, n, 0, bits, st);      // LSD radix: stable
    if (e == cudaSuccess) {
        k_mps_gather<<<mps_grid_blocks(n, 256, c->sm_count), 256, 0, st>>>(n, d_order, d_pos, d_pos_sorted);
        e = cudaGetLastError();
    }
    pool_release(tmp); pool_release(i1); pool_release(k2); pool_release(k1);
    if (e != cudaSuccess) { set_error("mgcg_mps_order: %s", cudaGetErrorString(e)); return -MGCG_ERR_CUDA; }
    return ncells;
}

// Row offsets (n + 1 ints) of the matrix of cell-ordered positions; returns the entry count, or a negative status
// (-MGCG_ERR_UNSUPPORTED when a row would exceed max_per_row, the reference's container limit).
extern "C" long long mgcg_mps_count(mgcg_ctx *c, int n, const double *d_pos_sorted, double re, int max_per_row, int *d_row_offsets)
{
    if (!c || n <= 0 || !(re > 0) || !d_pos_sorted || !d_row_offsets || max_per_row < 1) { set_error("mgcg_mps_count: bad argument"); return -MGCG_ERR_INVALID; }
    if (cudaSetDevice(c->device) != cudaSuccess) { set_error("mgcg_mps_count: cudaSetDevice failed"); return -MGCG_ERR_CUDA; }
    cudaStream_t st = c->stream;
    MpsGrid g;
    int s = mps_make_grid(c, n, d_pos_sorted, re, &g);
    if (s != MGCG_OK) return -s;
    int *cs = nullptr;
    s = mps_cell_start(c, n, g, d_pos_sorted, &cs);
    if (s != MGCG_OK) return -s;
    int *cnt = nullptr, *d_flag = nullptr;
    void *tmp = nullptr;
    size_t tb = 0;
    int hflag = 0, nnz = 0;
    cudaError_t e = pool_alloc((void **)&cnt, sizeof(int) * ((size_t)n + 1), st);
    if (e == cudaSuccess) e = pool_alloc((void **)&d_flag, sizeof(int), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_flag, 0, sizeof(int), st);
    if (e == cudaSuccess) {
        k_mps_rows<false><<<mps_grid_blocks((long long)n + 1, 128, c->sm_count), 128, 0, st>>>(n, g, d_pos_sorted, cs, 0.0, max_per_row, cnt,
                                                                                                nullptr, nullptr, nullptr, d_flag);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt, d_row_offsets, n + 1, st);
    if (e == cudaSuccess) e = pool_alloc(&tmp, tb, st);
    if (e == cudaSuccess) e = cub::DeviceScan::ExclusiveSum(tmp, tb, cnt, d_row_offsets, n + 1, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&hflag, d_flag, sizeof(int), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&nnz, d_row_offsets + n, sizeof(int), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    pool_release(tmp); pool_release(d_flag); pool_release(cnt); pool_release(cs);
    if (e != cudaSuccess) { set_error("mgcg_mps_count: %s", cudaGetErrorString(e)); return -MGCG_ERR_CUDA; }
    if (hflag) { set_error("mgcg_mps_count: a row has more than %d entries", max_per_row); return -MGCG_ERR_UNSUPPORTED; }
    if (nnz < 0) { set_error("mgcg_mps_count: entry count exceeds int32"); return -MGCG_ERR_UNSUPPORTED; }
    return nnz;
}

// Values and columns for the row offsets mgcg_mps_count produced.
extern "C" int mgcg_mps_fill(mgcg_ctx *c, int n, const double *d_pos_sorted, double re, double shift_abs, const int *d_row_offsets,
                             double *d_elements, int *d_column_indeces)
{
    if (!c || n <= 0 || !(re > 0) || !d_pos_sorted || !d_row_offsets || !d_elements || !d_column_indeces) { set_error("mgcg_mps_fill: bad argument"); return MGCG_ERR_INVALID; }
    MGCG_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    MpsGrid g;
    MGCG_CHECK(mps_make_grid(c, n, d_pos_sorted, re, &g));
    int *cs = nullptr;
    MGCG_CHECK(mps_cell_start(c, n, g, d_pos_sorted, &cs));
    k_mps_rows<true><<<mps_grid_blocks(n, 128, c->sm_count), 128, 0, st>>>(n, g, d_pos_sorted, cs, shift_abs, 0x7fffffff, nullptr, d_row_offsets,
                                                                           d_elements, d_column_indeces, nullptr);
    cudaError_t e = cudaGetLastError();
    pool_release(cs);
    if (e != cudaSuccess) { set_error("mgcg_mps_fill: %s", cudaGetErrorString(e)); return MGCG_ERR_CUDA; }
    return MGCG_OK;
}
