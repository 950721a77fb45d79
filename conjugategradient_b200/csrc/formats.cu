// formats.cu — the reference's other two input surfaces in front of the same solver (include/mgcg_formats.h):
// the ELL container of its OpenCL variant (Mgcg/HandmadeCL/MgcgCL/SparseMatrix.cs:13-47) and the four-call
// Write / Solve / Read / Iteration object of its ViennaCL variant (Mgcg/ViennaCL/Mgcg/ComputerGpu.hpp:68-103).
// Both end in mgcg_setup + mgcg_solve: the kernels stream CSR, so ELL is converted once, on the device, slot order kept.
#include "common.cuh"
#include "../../include/mgcg_formats.h"
#include <cub/cub.cuh>

namespace mgcg {

// counts -> a validated copy with a trailing zero (the scan's input); *bad counts rows outside [1, maxnz]; *total the sum
__global__ void k_ell_check(int n, int maxnz, const int *__restrict__ cnt, int *__restrict__ out, int *bad, unsigned long long *total)
{
    unsigned long long s = 0;
    int b = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= n; i += gridDim.x * blockDim.x) {
        int v = 0;
        if (i < n) {
            v = cnt[i];
            if (v < 1 || v > maxnz) { ++b; v = 0; }      // SparseMatrix.Clear leaves every row with its diagonal: count >= 1
        }
        out[i] = v;
        s += (unsigned long long)v;
    }
    for (int o = 16; o > 0; o >>= 1) { s += __shfl_down_sync(0xffffffffu, s, o); b += __shfl_down_sync(0xffffffffu, b, o); }
    if ((threadIdx.x & 31) == 0) {
        if (s) atomicAdd(total, s);
        if (b) atomicAdd(bad, b);
    }
}

// 8 lanes per row copy the used slots, in slot order, to the row's CSR segment
__global__ void __launch_bounds__(256) k_ell_fill(int n, int maxnz, const double *__restrict__ ev, const int *__restrict__ ec,
                                                  const int *__restrict__ rowptr, double *__restrict__ val, int *__restrict__ col)
{
    const int lane = threadIdx.x & 7;
    const long long groups = ((long long)gridDim.x * blockDim.x) >> 3;
    for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 3; i < n; i += groups) {
        const int o = rowptr[i], m = rowptr[i + 1] - o;
        const size_t base = (size_t)i * maxnz;
        for (int k = lane; k < m; k += 8) { val[o + k] = ev[base + k]; col[o + k] = ec[base + k]; }
    }
}

static int ell_grid(long long threads, int sm)
{
    long long g = (threads + 255) / 256;
    if (g > (long long)sm * 16) g = (long long)sm * 16;
    return (int)(g < 1 ? 1 : g);
}

// checked counts (+ trailing 0) into *d_cnt1 (pool block the caller releases); the total into *nnz
static int ell_counts(mgcg_ctx *c, int count, int maxnz, const int *d_counts, int **d_cnt1, long long *nnz)
{
    cudaStream_t st = c->stream;
    int *cnt1 = nullptr;
    unsigned long long *d_tot = nullptr;
    MGCG_CUDA(pool_alloc((void **)&cnt1, sizeof(int) * ((size_t)count + 1), st));
    MGCG_CUDA(pool_alloc((void **)&d_tot, 2 * sizeof(unsigned long long), st));      // [0] total, [1] (as int) bad rows
    MGCG_CUDA(cudaMemsetAsync(d_tot, 0, 2 * sizeof(unsigned long long), st));
    k_ell_check<<<ell_grid((long long)count + 1, c->sm_count), 256, 0, st>>>(count, maxnz, d_counts, cnt1, (int *)(d_tot + 1), d_tot);
    unsigned long long h[2] = {0, 0};
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(h, d_tot, sizeof(h), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    pool_release(d_tot);
    if (e != cudaSuccess) { pool_release(cnt1); set_error("ELL row counts: %s", cudaGetErrorString(e)); return MGCG_ERR_CUDA; }
    const int bad = (int)(h[1] & 0xffffffffu);
    if (bad) {
        pool_release(cnt1);
        set_error("ELL matrix: %d row(s) with a count outside [1, %d] (every row stores at least its diagonal, SparseMatrix.cs:52-63)", bad, maxnz);
        return MGCG_ERR_INVALID;
    }
    if (h[0] >= (1ull << 31)) { pool_release(cnt1); set_error("ELL matrix: %llu stored entries exceed int32", h[0]); return MGCG_ERR_UNSUPPORTED; }
    *d_cnt1 = cnt1; *nnz = (long long)h[0];
    return MGCG_OK;
}

static int ell_convert(mgcg_ctx *c, int count, int maxnz, const double *d_ev, const int *d_ec, int *d_cnt1,
                       double *d_val, int *d_rowptr, int *d_col)
{
    cudaStream_t st = c->stream;
    size_t tb = 0;
    MGCG_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, d_cnt1, d_rowptr, count + 1, st));
    void *tmp = nullptr;
    MGCG_CUDA(pool_alloc(&tmp, tb, st));
    cudaError_t e = cub::DeviceScan::ExclusiveSum(tmp, tb, d_cnt1, d_rowptr, count + 1, st);
    pool_release(tmp);
    if (e != cudaSuccess) { set_error("ELL scan: %s", cudaGetErrorString(e)); return MGCG_ERR_CUDA; }
    k_ell_fill<<<ell_grid((long long)count * 8, c->sm_count), 256, 0, st>>>(count, maxnz, d_ev, d_ec, d_rowptr, d_val, d_col);
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

// device buffers of the host-pointer ELL call and of the computer object
struct OwnedSystem {
    double *val = nullptr; int *rowptr = nullptr; int *col = nullptr; double *b = nullptr; double *x = nullptr;
    size_t cap_n = 0, cap_nnz = 0;
    void release()
    {
        if (val) cudaFree(val); if (rowptr) cudaFree(rowptr); if (col) cudaFree(col); if (b) cudaFree(b); if (x) cudaFree(x);
        val = nullptr; rowptr = nullptr; col = nullptr; b = nullptr; x = nullptr; cap_n = cap_nnz = 0;
    }
    int reserve(size_t n, size_t nnz)
    {
        if (cap_n >= n && cap_nnz >= nnz) return MGCG_OK;
        release();
        MGCG_CUDA(cudaMalloc((void **)&val, sizeof(double) * (nnz > 0 ? nnz : 1)));
        MGCG_CUDA(cudaMalloc((void **)&col, sizeof(int) * (nnz > 0 ? nnz : 1)));
        MGCG_CUDA(cudaMalloc((void **)&rowptr, sizeof(int) * (n + 1)));
        MGCG_CUDA(cudaMalloc((void **)&b, sizeof(double) * n));
        MGCG_CUDA(cudaMalloc((void **)&x, sizeof(double) * n));
        cap_n = n; cap_nnz = nnz;
        return MGCG_OK;
    }
};

}  // namespace mgcg

using namespace mgcg;

struct mgcg_computer {
    mgcg_ctx *ctx = nullptr;
    int n = 0;
    int nnz = 0;
    bool written = false;
    OwnedSystem sys;
    mgcg_options opt;
    mgcg_result last;
};

namespace mgcg {
void formats_free(mgcg_ctx *c)
{
    if (!c->ell_path) return;
    ((OwnedSystem *)c->ell_path)->release();
    delete (OwnedSystem *)c->ell_path;
    c->ell_path = nullptr;
}
}  // namespace mgcg

extern "C" {

// ---- (1) ELL ----------------------------------------------------------------------------------------------------
long long mgcg_ell_nnz(mgcg_ctx *c, int count, int max_nonzero_per_row, const int *d_nonzero_counts)
{
    if (!c || count <= 0 || max_nonzero_per_row < 1 || !d_nonzero_counts) { set_error("mgcg_ell_nnz: bad argument"); return -1; }
    if (cudaSetDevice(c->device) != cudaSuccess) { set_error("mgcg_ell_nnz: cudaSetDevice failed"); return -1; }
    int *cnt1 = nullptr;
    long long nnz = 0;
    if (ell_counts(c, count, max_nonzero_per_row, d_nonzero_counts, &cnt1, &nnz) != MGCG_OK) return -1;
    pool_release(cnt1);
    return nnz;
}

int mgcg_ell_to_csr(mgcg_ctx *c, int count, int max_nonzero_per_row,
                    const double *d_ell_elements, const int *d_ell_column_indeces, const int *d_nonzero_counts,
                    double *d_elements, int *d_row_offsets, int *d_column_indeces)
{
    if (!c || count <= 0 || max_nonzero_per_row < 1 || !d_ell_elements || !d_ell_column_indeces || !d_nonzero_counts ||
        !d_elements || !d_row_offsets || !d_column_indeces) {
        set_error("mgcg_ell_to_csr: bad argument");
        return MGCG_ERR_INVALID;
    }
    MGCG_CUDA(cudaSetDevice(c->device));
    int *cnt1 = nullptr;
    long long nnz = 0;
    MGCG_CHECK(ell_counts(c, count, max_nonzero_per_row, d_nonzero_counts, &cnt1, &nnz));
    int s = ell_convert(c, count, max_nonzero_per_row, d_ell_elements, d_ell_column_indeces, cnt1, d_elements, d_row_offsets, d_column_indeces);
    pool_release(cnt1);
    return s;
}

int mgcg_solve_host_ell(mgcg_ctx *c, int count, int max_nonzero_per_row,
                        const double *elements, const int *column_indeces, const int *nonzero_counts,
                        const double *b, double *x, const mgcg_options *opt, mgcg_result *res)
{
    if (!c || count <= 0 || max_nonzero_per_row < 1 || !elements || !column_indeces || !nonzero_counts || !b || !x) {
        set_error("mgcg_solve_host_ell: null or empty argument");
        return MGCG_ERR_INVALID;
    }
    MGCG_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const size_t slots = (size_t)count * (size_t)max_nonzero_per_row;
    // the ELL arrays travel as they are (ConjugateGradientSingleGpu.cs:214-216 does the same) and are dropped after the conversion
    double *d_ev = nullptr; int *d_ec = nullptr, *d_cnt = nullptr, *cnt1 = nullptr;
    MGCG_CUDA(pool_alloc((void **)&d_ev, sizeof(double) * slots, st));
    MGCG_CUDA(pool_alloc((void **)&d_ec, sizeof(int) * slots, st));
    MGCG_CUDA(pool_alloc((void **)&d_cnt, sizeof(int) * (size_t)count, st));
    int s = MGCG_OK;
    long long nnz = 0;
    cudaError_t e = cudaMemcpyAsync(d_cnt, nonzero_counts, sizeof(int) * (size_t)count, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_ev, elements, sizeof(double) * slots, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_ec, column_indeces, sizeof(int) * slots, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) { set_error("mgcg_solve_host_ell: upload: %s", cudaGetErrorString(e)); s = MGCG_ERR_CUDA; }
    if (s == MGCG_OK) s = ell_counts(c, count, max_nonzero_per_row, d_cnt, &cnt1, &nnz);
    OwnedSystem *h = nullptr;
    if (s == MGCG_OK) {
        if (!c->ell_path) c->ell_path = new OwnedSystem();
        h = (OwnedSystem *)c->ell_path;
        s = h->reserve((size_t)count, (size_t)nnz);
    }
    if (s == MGCG_OK) s = ell_convert(c, count, max_nonzero_per_row, d_ev, d_ec, cnt1, h->val, h->rowptr, h->col);
    pool_release(cnt1); pool_release(d_cnt); pool_release(d_ec); pool_release(d_ev);     // stream-ordered: reusable after the conversion
    if (s != MGCG_OK) return s;
    MGCG_CUDA(cudaMemcpyAsync(h->b, b, sizeof(double) * (size_t)count, cudaMemcpyHostToDevice, st));
    MGCG_CUDA(cudaMemcpyAsync(h->x, x, sizeof(double) * (size_t)count, cudaMemcpyHostToDevice, st));
    MGCG_CHECK(mgcg_setup(c, count, (int)nnz, h->val, h->rowptr, h->col, opt));
    s = mgcg_solve(c, h->b, h->x, res);
    MGCG_CUDA(cudaMemcpyAsync(x, h->x, sizeof(double) * (size_t)count, cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    return s;
}

// ---- (2) Write / Solve / Read / Iteration ---------------------------------------------------------------------------
mgcg_computer *mgcg_computer_create(int device, int n)
{
    if (n <= 0) { set_error("mgcg_computer_create: n = %d", n); return nullptr; }
    mgcg_ctx *ctx = mgcg_create(device);
    if (!ctx) return nullptr;
    mgcg_computer *k = new mgcg_computer();
    k->ctx = ctx; k->n = n;
    mgcg_options_default(&k->opt);
    memset(&k->last, 0, sizeof(k->last));
    return k;
}

void mgcg_computer_destroy(mgcg_computer *k)
{
    if (!k) return;
    if (k->ctx) { cudaSetDevice(k->ctx->device); if (k->ctx->stream) cudaStreamSynchronize(k->ctx->stream); }
    mgcg_destroy(k->ctx);
    k->sys.release();
    delete k;
}

int mgcg_computer_set_preconditioner(mgcg_computer *k, int precond, int nx, int ny, int nz)
{
    if (!k || (precond != MGCG_PRECOND_NONE && precond != MGCG_PRECOND_MG)) { set_error("mgcg_computer_set_preconditioner: bad argument"); return MGCG_ERR_INVALID; }
    k->opt.precond = precond; k->opt.nx = nx; k->opt.ny = ny; k->opt.nz = nz;
    return MGCG_OK;
}

int mgcg_computer_write(mgcg_computer *k, const double *elements, const unsigned int *row_offsets,
                        const unsigned int *column_indeces, int elements_count, const double *x, const double *b)
{
    if (!k || !elements || !row_offsets || !column_indeces || !x || !b || elements_count < 0) {
        set_error("mgcg_computer_write: null argument");
        return MGCG_ERR_INVALID;
    }
    // unsigned indices (ComputerGpu.hpp:47) below 2^31 have the bit pattern of the int32 the kernels read
    if (row_offsets[0] != 0u || row_offsets[k->n] != (unsigned)elements_count) {
        set_error("mgcg_computer_write: row_offsets[0] = %u, row_offsets[n] = %u, elements_count = %d", row_offsets[0], row_offsets[k->n], elements_count);
        return MGCG_ERR_INVALID;
    }
    mgcg_ctx *c = k->ctx;
    MGCG_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    k->written = false;
    MGCG_CHECK(k->sys.reserve((size_t)k->n, (size_t)elements_count));
    MGCG_CUDA(cudaMemcpyAsync(k->sys.val, elements, sizeof(double) * (size_t)elements_count, cudaMemcpyHostToDevice, st));
    MGCG_CUDA(cudaMemcpyAsync(k->sys.col, column_indeces, sizeof(int) * (size_t)elements_count, cudaMemcpyHostToDevice, st));
    MGCG_CUDA(cudaMemcpyAsync(k->sys.rowptr, row_offsets, sizeof(int) * ((size_t)k->n + 1), cudaMemcpyHostToDevice, st));
    MGCG_CUDA(cudaMemcpyAsync(k->sys.x, x, sizeof(double) * (size_t)k->n, cudaMemcpyHostToDevice, st));
    MGCG_CUDA(cudaMemcpyAsync(k->sys.b, b, sizeof(double) * (size_t)k->n, cudaMemcpyHostToDevice, st));
    MGCG_CUDA(cudaStreamSynchronize(st));      // the caller's arrays are free again (SetVector waits too, ComputerGpu.cpp:34)
    k->nnz = elements_count;
    k->written = true;
    return MGCG_OK;
}

int mgcg_computer_solve(mgcg_computer *k, double residual, int min_iteration, int max_iteration)
{
    if (!k || !k->written) { set_error("mgcg_computer_solve: nothing written"); return MGCG_ERR_INVALID; }
    mgcg_options o = k->opt;
    o.tolerance = residual;
    o.min_iteration = min_iteration;
    o.max_iteration = max_iteration > 0 ? max_iteration : (1 << 30);
    o.norm = MGCG_NORM_REL_L2;                    // rrNew / rr0 < residual^2, ComputerGpu.cpp:78
    o.min_strict = 1;                             // minIteration < iteration, ComputerGpu.cpp:78
    MGCG_CHECK(mgcg_setup(k->ctx, k->n, k->nnz, k->sys.val, k->sys.rowptr, k->sys.col, &o));
    return mgcg_solve(k->ctx, k->sys.b, k->sys.x, &k->last);
}

int mgcg_computer_read(mgcg_computer *k, double *x)
{
    if (!k || !x || !k->written) { set_error("mgcg_computer_read: nothing written"); return MGCG_ERR_INVALID; }
    MGCG_CUDA(cudaSetDevice(k->ctx->device));
    MGCG_CUDA(cudaMemcpyAsync(x, k->sys.x, sizeof(double) * (size_t)k->n, cudaMemcpyDeviceToHost, k->ctx->stream));
    MGCG_CUDA(cudaStreamSynchronize(k->ctx->stream));
    return MGCG_OK;
}

int mgcg_computer_iteration(const mgcg_computer *k) { return k ? k->last.iterations_done : -1; }

int mgcg_computer_result(const mgcg_computer *k, mgcg_result *res)
{
    if (!k || !res) { set_error("mgcg_computer_result: null argument"); return MGCG_ERR_INVALID; }
    *res = k->last;
    return MGCG_OK;
}

}  // extern "C"
