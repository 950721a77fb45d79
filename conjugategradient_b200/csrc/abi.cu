// abi.cu — the mgcg_* C ABI (include/mgcg.h): context lifetime, setup/solve on device and host pointers,
// individually exported kernels, the reference partition rule.
#include "mg.cuh"
#include <chrono>
#include <math.h>
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>

namespace mgcg {

static thread_local char g_err[1024] = "";

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
const char *get_error() { return g_err; }

bool pdl_enabled()
{
    // opt-in: measured SLOWER inside the loop's WHILE graph (256^3: 34.0 against 28.6 ms, 128^3: 5.86 against 5.38 ms,
    // profiles/r02_pdl_slab_owner.txt) — programmatic edges cost the graph ~7 us per kernel node, more than the overlapped prologue saves
    static const bool on = [] { const char *e = getenv("MGCG_PDL"); return e && e[0] == '1'; }();
    return on;
}

int ctx_ensure_work(mgcg_ctx *c, int n)
{
    if (c->work_n >= n && c->r) return MGCG_OK;
    if (c->r) cudaFree(c->r);
    if (c->p) cudaFree(c->p);
    if (c->Ap) cudaFree(c->Ap);
    if (c->s) cudaFree(c->s);
    c->r = c->p = c->Ap = c->s = nullptr;
    size_t bytes = sizeof(double) * (size_t)(n > 0 ? n : 1);
    MGCG_CUDA(cudaMalloc((void **)&c->r, bytes));
    MGCG_CUDA(cudaMalloc((void **)&c->p, bytes));
    MGCG_CUDA(cudaMalloc((void **)&c->Ap, bytes));
    c->work_n = n;
    if (c->graph_exec) { cudaGraphExecDestroy(c->graph_exec); c->graph_exec = nullptr; }
    if (c->graph) { cudaGraphDestroy(c->graph); c->graph = nullptr; }
    return MGCG_OK;
}

struct HostPath {   // device copies owned by mgcg_solve_host
    double *val = nullptr; int *rowptr = nullptr; int *col = nullptr; double *b = nullptr; double *x = nullptr;
    size_t cap_n = 0, cap_nnz = 0;
    cudaStream_t copy = nullptr;          // b and x0 travel on this stream while the hierarchy is built on the solve stream
    cudaEvent_t ev_matrix = nullptr, ev_vectors = nullptr;
    void release()
    {
        if (val) cudaFree(val); if (rowptr) cudaFree(rowptr); if (col) cudaFree(col); if (b) cudaFree(b); if (x) cudaFree(x);
        val = nullptr; rowptr = nullptr; col = nullptr; b = nullptr; x = nullptr; cap_n = cap_nnz = 0;
    }
    void release_all()
    {
        release();
        if (copy) cudaStreamDestroy(copy);
        if (ev_matrix) cudaEventDestroy(ev_matrix);
        if (ev_vectors) cudaEventDestroy(ev_vectors);
        copy = nullptr; ev_matrix = nullptr; ev_vectors = nullptr;
    }
};

}  // namespace mgcg

using namespace mgcg;

extern "C" {

int mgcg_version(void) { return MGCG_VERSION; }
const char *mgcg_last_error(void) { return get_error(); }

void mgcg_options_default(mgcg_options *o)
{
    memset(o, 0, sizeof(*o));
    o->tolerance = 1e-8;
    o->min_iteration = 0;
    o->max_iteration = 1 << 30;
    o->norm = MGCG_NORM_ABS_L2;
    o->min_strict = 0;
    o->precond = MGCG_PRECOND_NONE;
    o->mg_max_levels = 12;
    o->mg_coarse_size = 1400;         /* 11^3 = 1331 rows are still solved densely (blocked Gauss-Jordan inverse): one level fewer in the cycle than with 512 */
    o->mg_pre_sweeps = 1;
    o->mg_post_sweeps = 1;
    o->mg_omega = 0.85;   /* ~6/7, the optimal Jacobi damping of the 3-D 7-point stencil: one iteration fewer than 0.8 at 256^3 */
    o->mg_smoother = MGCG_SMOOTHER_JACOBI;
    o->mg_smoothed = 1;
    o->mg_aggregate_size = 3;          /* 3x3x3 on the finest level, 2x2x2 below: fewest total bytes per digit of residual */
    o->mg_aggregate_size_coarse = 2;
    o->dist_replicate_below = 1 << 18;
    o->mg_prolong_omega = 2.0 / 3.0;
    o->spmv_kernel = MGCG_SPMV_AUTO;
    o->check_interval = 8;
    o->use_graph = 2;
    o->profile = 0;
    o->mg_aggregation = MGCG_AGG_AUTO;
    o->mg_strength_theta = 0.5;
    o->mg_cheby_degree = 2;
    o->mg_cheby_ratio = 6.0;
    o->cg_variant = MGCG_CG_CLASSIC;
}

mgcg_ctx *mgcg_create(int device)
{
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0) {
        set_error("mgcg_create: no CUDA device (%s) - this library has no CPU fallback", cudaGetErrorString(e));
        return nullptr;
    }
    if (device < 0 || device >= count) { set_error("mgcg_create: device %d out of range [0,%d)", device, count); return nullptr; }
    if (cudaSetDevice(device) != cudaSuccess) { set_error("mgcg_create: cudaSetDevice(%d) failed", device); return nullptr; }
    mgcg_ctx *c = new mgcg_ctx();
    c->device = device;
    cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device);
    if (c->sm_count <= 0) c->sm_count = 148;
    bool ok = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) == cudaSuccess;
    c->own_stream = true;
    ok = ok && cudaMalloc((void **)&c->d_scalars, sizeof(Scalars)) == cudaSuccess;
    ok = ok && cudaMallocHost((void **)&c->h_scalars, 2 * sizeof(Scalars)) == cudaSuccess;
    ok = ok && cudaMalloc((void **)&c->ws.partial, sizeof(double) * 2 * kMaxPartials) == cudaSuccess;
    ok = ok && cudaMalloc((void **)&c->ws.ticket, sizeof(unsigned int)) == cudaSuccess;
    ok = ok && cudaMalloc((void **)&c->d_tmp_scalar, sizeof(double) * 4) == cudaSuccess;
    ok = ok && cudaMallocHost((void **)&c->h_tmp_scalar, sizeof(double) * 4) == cudaSuccess;
    ok = ok && cudaMalloc((void **)&c->tail_barrier, 2 * sizeof(unsigned int)) == cudaSuccess;
    ok = ok && cudaMemset(c->tail_barrier, 0, 2 * sizeof(unsigned int)) == cudaSuccess;
    ok = ok && cudaMemset(c->ws.ticket, 0, sizeof(unsigned int)) == cudaSuccess;
    ok = ok && cudaMemset(c->d_scalars, 0, sizeof(Scalars)) == cudaSuccess;
    if (!ok) { set_error("mgcg_create: allocation failed: %s", cudaGetErrorString(cudaGetLastError())); mgcg_destroy(c); return nullptr; }
    {   // keep freed hierarchy-build blocks in the device pool instead of returning them to the driver at every sync
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
            unsigned long long keep = ~0ull;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
    }
    mgcg_options_default(&c->opt);
    return c;
}

void mgcg_destroy(mgcg_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    dist_free(c);
    mg_free(c);
    spmv_code_free(&c->code0);
    if (!c->child) pool_trim_stream(c->device, c->stream);
    if (c->graph_exec) cudaGraphExecDestroy(c->graph_exec);
    if (c->graph) cudaGraphDestroy(c->graph);
    if (c->host_path) { ((HostPath *)c->host_path)->release_all(); delete (HostPath *)c->host_path; }
    formats_free(c);
    dist_host_path_free(c);
    if (c->r) cudaFree(c->r);
    if (c->p) cudaFree(c->p);
    if (c->Ap) cudaFree(c->Ap);
    if (c->s) cudaFree(c->s);
    if (c->d_scalars) cudaFree(c->d_scalars);
    if (c->h_scalars) cudaFreeHost(c->h_scalars);
    if (c->ws.partial) cudaFree(c->ws.partial);
    if (c->ws.ticket) cudaFree(c->ws.ticket);
    if (c->tail_barrier) cudaFree(c->tail_barrier);
    if (c->d_tmp_scalar) cudaFree(c->d_tmp_scalar);
    if (c->h_tmp_scalar) cudaFreeHost(c->h_tmp_scalar);
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int mgcg_set_stream(mgcg_ctx *c, void *cuda_stream)
{
    if (!c) { set_error("mgcg_set_stream: null context"); return MGCG_ERR_INVALID; }
    MGCG_CUDA(cudaSetDevice(c->device));
    if (c->stream) MGCG_CUDA(cudaStreamSynchronize(c->stream));
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    c->stream = (cudaStream_t)cuda_stream;
    c->own_stream = false;
    if (c->graph_exec) { cudaGraphExecDestroy(c->graph_exec); c->graph_exec = nullptr; }
    if (c->graph) { cudaGraphDestroy(c->graph); c->graph = nullptr; }
    return MGCG_OK;
}

int mgcg_setup(mgcg_ctx *c, int count, int elements_count,
               const double *d_elements, const int *d_row_offsets, const int *d_column_indeces, const mgcg_options *opt)
{
    if (!c) { set_error("mgcg_setup: null context"); return MGCG_ERR_INVALID; }
    if (count <= 0 || elements_count < 0 || !d_elements || !d_row_offsets || !d_column_indeces) {
        set_error("mgcg_setup: empty or null matrix (count=%d, elements=%d)", count, elements_count);
        return MGCG_ERR_INVALID;
    }
    MGCG_CUDA(cudaSetDevice(c->device));
    if (opt) c->opt = *opt;
    c->A.rows = count; c->A.cols = count; c->A.nnz = elements_count;
    c->A.val = d_elements; c->A.rowptr = d_row_offsets; c->A.col = d_column_indeces;
    c->planA = spmv_make_plan(c, count, elements_count, c->opt.spmv_kernel);
    if (c->graph_exec) { cudaGraphExecDestroy(c->graph_exec); c->graph_exec = nullptr; }
    if (c->graph) { cudaGraphDestroy(c->graph); c->graph = nullptr; }
    c->is_setup = false;
    c->setup_ms = 0.f;
    MGCG_CHECK(ctx_ensure_work(c, count));
    if (c->opt.precond == MGCG_PRECOND_MG) {
        // the block cache's next slab: what a hierarchy build of this size takes (operators, transients, sort buffers, level vectors)
        pool_reserve((size_t)(2.2 * 12.0 * (double)elements_count + 200.0 * (double)count) + ((size_t)64 << 20));
        cudaEvent_t e0, e1;
        MGCG_CUDA(cudaEventCreate(&e0)); MGCG_CUDA(cudaEventCreate(&e1));
        MGCG_CUDA(cudaEventRecord(e0, c->stream));
        int s = spmv_code_build(c, &c->code0, &c->A, &c->planA);       // coded columns of the fine operator (timed with the set-up)
        if (s == MGCG_OK) s = mg_build(c);
        if (s != MGCG_OK) { mg_free(c); cudaEventDestroy(e0); cudaEventDestroy(e1); return s; }
        MGCG_CUDA(cudaEventRecord(e1, c->stream));
        MGCG_CUDA(cudaEventSynchronize(e1));
        MGCG_CUDA(cudaEventElapsedTime(&c->setup_ms, e0, e1));
        cudaEventDestroy(e0); cudaEventDestroy(e1);
    } else {
        mg_free(c);
        MGCG_CHECK(spmv_code_build(c, &c->code0, &c->A, &c->planA));
    }
    c->is_setup = true;
    return MGCG_OK;
}

int mgcg_solve(mgcg_ctx *c, const double *d_b, double *d_x, mgcg_result *res)
{
    if (!c || !d_b || !d_x) { set_error("mgcg_solve: null argument"); return MGCG_ERR_INVALID; }
    MGCG_CUDA(cudaSetDevice(c->device));
    if (res) memset(res, 0, sizeof(*res));
    int s = cg_solve(c, d_b, d_x, res);
    if (res) {
        res->status = s;
        res->operator_complexity = 1.0;
        if (c->mg) res->operator_complexity = c->mg->op_complexity;
    }
    return s;
}

int mgcg_solve_host(mgcg_ctx *c, int count, int elements_count,
                    const double *elements, const int *row_offsets, const int *column_indeces,
                    const double *b, double *x, const mgcg_options *opt, mgcg_result *res)
{
    if (!c || !elements || !row_offsets || !column_indeces || !b || !x || count <= 0) {
        set_error("mgcg_solve_host: null or empty argument");
        return MGCG_ERR_INVALID;
    }
    MGCG_CUDA(cudaSetDevice(c->device));
    if (!c->host_path) c->host_path = new HostPath();
    HostPath *h = (HostPath *)c->host_path;
    if (h->cap_n < (size_t)count || h->cap_nnz < (size_t)elements_count) {
        h->release();
        MGCG_CUDA(cudaMalloc((void **)&h->val, sizeof(double) * (size_t)(elements_count > 0 ? elements_count : 1)));
        MGCG_CUDA(cudaMalloc((void **)&h->col, sizeof(int) * (size_t)(elements_count > 0 ? elements_count : 1)));
        MGCG_CUDA(cudaMalloc((void **)&h->rowptr, sizeof(int) * ((size_t)count + 1)));
        MGCG_CUDA(cudaMalloc((void **)&h->b, sizeof(double) * (size_t)count));
        MGCG_CUDA(cudaMalloc((void **)&h->x, sizeof(double) * (size_t)count));
        h->cap_n = count; h->cap_nnz = elements_count;
    }
    cudaStream_t st = c->stream;
    if (!h->copy) {
        MGCG_CUDA(cudaStreamCreateWithFlags(&h->copy, cudaStreamNonBlocking));
        MGCG_CUDA(cudaEventCreateWithFlags(&h->ev_matrix, cudaEventDisableTiming));
        MGCG_CUDA(cudaEventCreateWithFlags(&h->ev_vectors, cudaEventDisableTiming));
    }
    // the matrix first, on the solve stream (the hierarchy build needs all of it); b and x0 follow on a second stream once the
    // matrix has left the host, so that their 16 N bytes cross PCIe while the hierarchy is being built
    const bool trace = getenv("MGCG_TRACE") != nullptr;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto ms_since = [](std::chrono::steady_clock::time_point a) { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - a).count(); };
    auto t0 = now();
    MGCG_CUDA(cudaMemcpyAsync(h->rowptr, row_offsets, sizeof(int) * ((size_t)count + 1), cudaMemcpyHostToDevice, st));
    MGCG_CUDA(cudaMemcpyAsync(h->col, column_indeces, sizeof(int) * (size_t)elements_count, cudaMemcpyHostToDevice, st));
    MGCG_CUDA(cudaMemcpyAsync(h->val, elements, sizeof(double) * (size_t)elements_count, cudaMemcpyHostToDevice, st));
    MGCG_CUDA(cudaEventRecord(h->ev_matrix, st));
    MGCG_CUDA(cudaStreamWaitEvent(h->copy, h->ev_matrix, 0));
    MGCG_CUDA(cudaMemcpyAsync(h->b, b, sizeof(double) * (size_t)count, cudaMemcpyHostToDevice, h->copy));
    MGCG_CUDA(cudaMemcpyAsync(h->x, x, sizeof(double) * (size_t)count, cudaMemcpyHostToDevice, h->copy));
    MGCG_CUDA(cudaEventRecord(h->ev_vectors, h->copy));
    if (trace) { cudaEventSynchronize(h->ev_matrix); fprintf(stderr, "[mgcg host] matrix on the device   %8.3f ms\n", ms_since(t0)); }
    int s = mgcg_setup(c, count, elements_count, h->val, h->rowptr, h->col, opt);
    if (trace) fprintf(stderr, "[mgcg host] hierarchy built        %8.3f ms (vectors %s)\n", ms_since(t0), cudaEventQuery(h->ev_vectors) == cudaSuccess ? "arrived meanwhile" : "still in flight");
    // the solve (and any early return) must not overtake the vector upload
    MGCG_CUDA(cudaStreamWaitEvent(st, h->ev_vectors, 0));
    if (s != MGCG_OK) { cudaStreamSynchronize(h->copy); return s; }
    s = mgcg_solve(c, h->b, h->x, res);
    if (trace) fprintf(stderr, "[mgcg host] solved                 %8.3f ms\n", ms_since(t0));
    MGCG_CUDA(cudaMemcpyAsync(x, h->x, sizeof(double) * (size_t)count, cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    if (trace) fprintf(stderr, "[mgcg host] x on the host          %8.3f ms\n", ms_since(t0));
    return s;
}

// ---- individually exported kernels ----------------------------------------------------------------------
static int fetch_scalar(mgcg_ctx *c, const double *d, double *host_out)
{
    if (!host_out) return MGCG_OK;
    MGCG_CUDA(cudaMemcpyAsync(c->h_tmp_scalar, d, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    MGCG_CUDA(cudaStreamSynchronize(c->stream));
    *host_out = c->h_tmp_scalar[0];
    return MGCG_OK;
}

// the arrays of the matrix the context was set up with (then the launch may use what the set-up derived from them)
static bool same_as_setup(const mgcg_ctx *c, const Csr &A)
{
    return c->is_setup && A.rows == c->A.rows && A.cols == c->A.cols && A.nnz == c->A.nnz &&
           A.val == c->A.val && A.rowptr == c->A.rowptr && A.col == c->A.col;
}

int mgcg_matrix_codes(mgcg_ctx *c, int *ndict, int *dict, unsigned char *codes)
{
    if (!c || !c->is_setup) { set_error("mgcg_matrix_codes: context not set up"); return MGCG_ERR_INVALID; }
    MGCG_CUDA(cudaSetDevice(c->device));
    const bool coded = c->A.code && c->A.dict;
    if (ndict) *ndict = coded ? c->A.ndict : 0;
    if (!coded) return MGCG_OK;
    MGCG_CUDA(cudaStreamSynchronize(c->stream));
    if (dict) MGCG_CUDA(cudaMemcpy(dict, c->A.dict, sizeof(int) * 256, cudaMemcpyDeviceToHost));
    if (codes) MGCG_CUDA(cudaMemcpy(codes, c->A.code, (size_t)c->A.nnz, cudaMemcpyDeviceToHost));
    return MGCG_OK;
}

int mgcg_spmv(mgcg_ctx *c, int rows, int cols, int elements_count,
              const double *d_elements, const int *d_row_offsets, const int *d_column_indeces,
              const double *d_x, double *d_y, double alpha, double beta, int kernel)
{
    if (!c) { set_error("mgcg_spmv: null context"); return MGCG_ERR_INVALID; }
    MGCG_CUDA(cudaSetDevice(c->device));
    SpmvArgs a;
    a.A.rows = rows; a.A.cols = cols; a.A.nnz = elements_count; a.A.val = d_elements; a.A.rowptr = d_row_offsets; a.A.col = d_column_indeces;
    a.x = d_x; a.y = d_y; a.alpha = alpha; a.beta = beta;
    SpmvPlan p = spmv_make_plan(c, rows, elements_count, kernel);
    if (same_as_setup(c, a.A) && kernel == c->opt.spmv_kernel) { a.A = c->A; p = c->planA; dist_x_range(c, &a.x_lo, &a.x_hi); }   // the set-up matrix: its coded columns
    const bool plain = (alpha == 1.0 && beta == 0.0);
    return spmv_launch(c, p, plain ? EPI_PLAIN : EPI_AXPBY, a);
}

int mgcg_spmv_dot(mgcg_ctx *c, int rows, int cols, int elements_count,
                  const double *d_elements, const int *d_row_offsets, const int *d_column_indeces,
                  const double *d_x, double *d_y, const double *d_w, double *d_dot, int kernel)
{
    if (!c) { set_error("mgcg_spmv_dot: null context"); return MGCG_ERR_INVALID; }
    MGCG_CUDA(cudaSetDevice(c->device));
    SpmvArgs a;
    a.A.rows = rows; a.A.cols = cols; a.A.nnz = elements_count; a.A.val = d_elements; a.A.rowptr = d_row_offsets; a.A.col = d_column_indeces;
    a.x = d_x; a.y = d_y; a.w = d_w; a.fin = FIN_STORE; a.out = d_dot;
    SpmvPlan p = spmv_make_plan(c, rows, elements_count, kernel);
    if (same_as_setup(c, a.A) && kernel == c->opt.spmv_kernel) { a.A = c->A; p = c->planA; dist_x_range(c, &a.x_lo, &a.x_hi); }
    return spmv_launch(c, p, EPI_DOT, a);
}

int mgcg_dot(mgcg_ctx *c, int n, const double *d_a, const double *d_b, double *d_out, double *host_out)
{
    if (!c) { set_error("mgcg_dot: null context"); return MGCG_ERR_INVALID; }
    MGCG_CUDA(cudaSetDevice(c->device));
    double *out = d_out ? d_out : c->d_tmp_scalar;
    MGCG_CHECK(vec_dot(c, n, d_a, d_b, FIN_STORE, out));
    return fetch_scalar(c, out, host_out);
}

int mgcg_max_abs(mgcg_ctx *c, int n, const double *d_a, double *d_out, double *host_out)
{
    if (!c) { set_error("mgcg_max_abs: null context"); return MGCG_ERR_INVALID; }
    MGCG_CUDA(cudaSetDevice(c->device));
    double *out = d_out ? d_out : c->d_tmp_scalar;
    MGCG_CHECK(vec_max_abs(c, n, d_a, out));
    return fetch_scalar(c, out, host_out);
}

int mgcg_axpy(mgcg_ctx *c, int n, double alpha, const double *d_x, double *d_y)
{
    if (!c) return MGCG_ERR_INVALID;
    MGCG_CUDA(cudaSetDevice(c->device));
    return vec_axpy(c, n, alpha, d_x, d_y);
}
int mgcg_scal(mgcg_ctx *c, int n, double alpha, double *d_x)
{
    if (!c) return MGCG_ERR_INVALID;
    MGCG_CUDA(cudaSetDevice(c->device));
    return vec_scal(c, n, alpha, d_x);
}
int mgcg_copy(mgcg_ctx *c, int n, const double *d_x, double *d_y)
{
    if (!c) return MGCG_ERR_INVALID;
    MGCG_CUDA(cudaSetDevice(c->device));
    return vec_copy(c, n, d_x, d_y);
}

int mgcg_update_xr(mgcg_ctx *c, int n, double alpha, const double *d_p, const double *d_Ap,
                   double *d_x, double *d_r, double *d_rr, double *host_rr)
{
    if (!c) return MGCG_ERR_INVALID;
    MGCG_CUDA(cudaSetDevice(c->device));
    double *out = d_rr ? d_rr : c->d_tmp_scalar;
    MGCG_CHECK(cg_update_xr(c, n, d_p, d_Ap, d_x, d_r, FIN_STORE, true, alpha, out));
    return fetch_scalar(c, out, host_rr);
}

int mgcg_update_p(mgcg_ctx *c, int n, double beta, const double *d_r, double *d_p)
{
    if (!c) return MGCG_ERR_INVALID;
    MGCG_CUDA(cudaSetDevice(c->device));
    return cg_update_p(c, n, d_r, d_p, true, beta);
}

// ConjugateGradientParallelGpu.cs:271-277 — bit-exact integer rule
int mgcg_partition_offsets(int count, int device_count, int *offsets)
{
    if (device_count <= 0 || !offsets || count < 0) { set_error("mgcg_partition_offsets: bad argument"); return MGCG_ERR_INVALID; }
    offsets[0] = 0;
    for (int i = 1; i < device_count; i++) offsets[i] = offsets[i - 1] + (int)floor((double)count / device_count);
    offsets[device_count] = count;
    return MGCG_OK;
}

}  // extern "C"
