// compat.cu — the reference's 32 MgcgGpu.dll export names (include/mgcg_compat.h) on top of the native kernels.
// Handles are per-device contexts of this library; vectors are plain device allocations.
#include "common.cuh"
#include "../../include/mgcg_compat.h"
#include <mutex>

struct MgcgVector { double *d; int size; };
struct MgcgVectorInt { int *d; int size; };
struct MgcgBlas { mgcg_ctx *ctx; };
struct MgcgSparse { mgcg_ctx *ctx; };
struct MgcgMatDescr { int unused; };

using namespace mgcg;

namespace {

std::mutex g_mu;
mgcg_ctx *g_ctx[64] = {nullptr};
int g_precond = MGCG_PRECOND_NONE, g_nx = 0, g_ny = 0, g_nz = 0;

[[noreturn]] void die(const char *where)
{
    // The reference ABI has no error channel (void returns, statuses dropped); a failed call must not pass silently.
    fprintf(stderr, "MgcgGpu: %s failed: %s\n", where, mgcg_last_error());
    abort();
}

mgcg_ctx *current_ctx()
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { set_error("no CUDA device available (no CPU fallback)"); die("cudaGetDevice"); }
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_ctx[dev]) {
        g_ctx[dev] = mgcg_create(dev);
        if (!g_ctx[dev]) die("mgcg_create");
    }
    return g_ctx[dev];
}

void ck(int s, const char *where) { if (s != MGCG_OK) die(where); }
void ckc(cudaError_t e, const char *where) { if (e != cudaSuccess) { set_error("%s", cudaGetErrorString(e)); die(where); } }

__global__ void k_rebase(int n, int *a, int off)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) a[i] -= off;
}
__global__ void k_minmax_int(int n, const int *a, int *out)   // out[0]=min, out[1]=max (pre-initialised)
{
    int mn = 0x7fffffff, mx = -0x7fffffff - 1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) { int v = a[i]; mn = min(mn, v); mx = max(mx, v); }
    for (int o = 16; o > 0; o >>= 1) { mn = min(mn, __shfl_down_sync(0xffffffffu, mn, o)); mx = max(mx, __shfl_down_sync(0xffffffffu, mx, o)); }
    if ((threadIdx.x & 31) == 0) { atomicMin(&out[0], mn); atomicMax(&out[1], mx); }
}

double fetch(mgcg_ctx *c, const double *d)
{
    ckc(cudaMemcpyAsync(c->h_tmp_scalar, d, sizeof(double), cudaMemcpyDeviceToHost, c->stream), "cudaMemcpyAsync");
    ckc(cudaStreamSynchronize(c->stream), "cudaStreamSynchronize");
    return c->h_tmp_scalar[0];
}

}  // namespace

extern "C" {

// ---- Runtime.cu:7-62
int GetDeviceCount(void) { int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess) return 0; return n; }
void SetDevice(int deviceID) { ckc(cudaSetDevice(deviceID), "SetDevice"); }
MgcgBlas *CreateBlas(void) { MgcgBlas *h = new MgcgBlas(); h->ctx = current_ctx(); return h; }
void DestroyBlas(MgcgBlas *h) { delete h; }
MgcgSparse *CreateSparse(void) { MgcgSparse *h = new MgcgSparse(); h->ctx = current_ctx(); return h; }
void DestroySparse(MgcgSparse *h) { delete h; }
MgcgMatDescr *CreateMatDescr(void) { return new MgcgMatDescr(); }
void DestroyMatDescr(MgcgMatDescr *d) { delete d; }

// ---- Vector_Double.cu:7-45
MgcgVector *Create_Double(int size)
{
    current_ctx();
    MgcgVector *v = new MgcgVector();
    v->size = size;
    ckc(cudaMalloc((void **)&v->d, sizeof(double) * (size_t)(size > 0 ? size : 1)), "Create_Double");
    ckc(cudaMemset(v->d, 0, sizeof(double) * (size_t)(size > 0 ? size : 1)), "Create_Double");   // device_vector(n) value-initialises
    return v;
}
void CopyToArray_Double(const MgcgVector *source, double destination[], int count, int sourceOffset, int destinationOffset)
{
    mgcg_ctx *c = current_ctx();
    ckc(cudaStreamSynchronize(c->stream), "CopyToArray_Double");
    ckc(cudaMemcpy(destination + destinationOffset, source->d + sourceOffset, sizeof(double) * (size_t)count, cudaMemcpyDeviceToHost), "CopyToArray_Double");
}
void CopyFromArray_Double(MgcgVector *destination, const double source[], int count, int sourceOffset, int destinationOffset)
{
    mgcg_ctx *c = current_ctx();
    ckc(cudaStreamSynchronize(c->stream), "CopyFromArray_Double");
    ckc(cudaMemcpy(destination->d + destinationOffset, source + sourceOffset, sizeof(double) * (size_t)count, cudaMemcpyHostToDevice), "CopyFromArray_Double");
}
void Delete_Double(MgcgVector *vec) { if (vec) { cudaFree(vec->d); delete vec; } }
double *ToRawPtr_Double(MgcgVector *vec) { return vec->d; }
void CopyFromDevice_Double(const double *source, double *destination, int count, int sourceOffset, int destinationOffset)
{
    mgcg_ctx *c = current_ctx();
    ckc(cudaMemcpyAsync(destination + destinationOffset, source + sourceOffset, sizeof(double) * (size_t)count, cudaMemcpyDeviceToDevice, c->stream), "CopyFromDevice_Double");
    ckc(cudaStreamSynchronize(c->stream), "CopyFromDevice_Double");
}

// ---- Vector_Int.cu:9-40
MgcgVectorInt *Create_Int(int size)
{
    current_ctx();
    MgcgVectorInt *v = new MgcgVectorInt();
    v->size = size;
    ckc(cudaMalloc((void **)&v->d, sizeof(int) * (size_t)(size > 0 ? size : 1)), "Create_Int");
    ckc(cudaMemset(v->d, 0, sizeof(int) * (size_t)(size > 0 ? size : 1)), "Create_Int");
    return v;
}
void CopyToArray_Int(const MgcgVectorInt *source, int destination[], int count, int sourceOffset, int destinationOffset)
{
    mgcg_ctx *c = current_ctx();
    ckc(cudaStreamSynchronize(c->stream), "CopyToArray_Int");
    ckc(cudaMemcpy(destination + destinationOffset, source->d + sourceOffset, sizeof(int) * (size_t)count, cudaMemcpyDeviceToHost), "CopyToArray_Int");
}
void CopyFromArray_Int(MgcgVectorInt *destination, int source[], int count, int sourceOffset, int destinationOffset)
{
    mgcg_ctx *c = current_ctx();
    ckc(cudaStreamSynchronize(c->stream), "CopyFromArray_Int");
    ckc(cudaMemcpy(destination->d + destinationOffset, source + sourceOffset, sizeof(int) * (size_t)count, cudaMemcpyHostToDevice), "CopyFromArray_Int");
}
void Delete_Int(MgcgVectorInt *vec) { if (vec) { cudaFree(vec->d); delete vec; } }
int *ToRawPtr_Int(MgcgVectorInt *vec) { return vec->d; }

// ---- Mgcg.cu:10-54
void CsrMV(MgcgSparse *cusparse, MgcgMatDescr *, double *y, const double *elements, const int *rowOffsets, const int *columnIndeces,
           const double *x, int elementsCount, int rowCount, int columnCount, double alpha, double beta)
{
    mgcg_ctx *c = cusparse ? cusparse->ctx : current_ctx();
    ck(mgcg_spmv(c, rowCount, columnCount, elementsCount, elements, rowOffsets, columnIndeces, x, y, alpha, beta, MGCG_SPMV_AUTO), "CsrMV");
}
void Axpy(MgcgBlas *cublas, double *y, const double *x, int count, double alpha)
{
    mgcg_ctx *c = cublas ? cublas->ctx : current_ctx();
    ck(mgcg_axpy(c, count, alpha, x, y), "Axpy");
}
double Dot(MgcgBlas *cublas, double *y, const double *x, int count)
{
    mgcg_ctx *c = cublas ? cublas->ctx : current_ctx();
    double r = 0.0;
    ck(mgcg_dot(c, count, x, y, nullptr, &r), "Dot");
    return r;
}
void Scal(MgcgBlas *cublas, double *x, double alpha, int count)
{
    mgcg_ctx *c = cublas ? cublas->ctx : current_ctx();
    ck(mgcg_scal(c, count, alpha, x), "Scal");
}
void Copy(MgcgBlas *cublas, double *y, const double *x, int count, int yOffset, int xOffset)
{
    mgcg_ctx *c = cublas ? cublas->ctx : current_ctx();
    ck(mgcg_copy(c, count, x + xOffset, y + yOffset), "Copy");
}

// ---- Mgcg.cu:57-85
void Initialize(const double elements[], const int rowOffsets[], const int columnIndeces[], const double x[], const double b[],
                MgcgVector *elementsVector, MgcgVectorInt *rowOffsetsVector, MgcgVectorInt *columnIndecesVector,
                MgcgVector *xVector, MgcgVector *bVector, MgcgVector *pVector, int *minJ, int *maxJ, int count,
                int countForDevice, int offsetForDevice, int elementCountForDevice, int elementOffsetForDevice)
{
    (void)count;
    mgcg_ctx *c = current_ctx();
    cudaStream_t st = c->stream;
    ckc(cudaMemcpyAsync(elementsVector->d, elements + elementOffsetForDevice, sizeof(double) * (size_t)elementCountForDevice, cudaMemcpyHostToDevice, st), "Initialize");
    ckc(cudaMemcpyAsync(columnIndecesVector->d, columnIndeces + elementOffsetForDevice, sizeof(int) * (size_t)elementCountForDevice, cudaMemcpyHostToDevice, st), "Initialize");
    ckc(cudaMemcpyAsync(rowOffsetsVector->d, rowOffsets + offsetForDevice, sizeof(int) * ((size_t)countForDevice + 1), cudaMemcpyHostToDevice, st), "Initialize");
    k_rebase<<<div_up(countForDevice + 1, 256), 256, 0, st>>>(countForDevice + 1, rowOffsetsVector->d, elementOffsetForDevice);   // Mgcg.cu:73
    ckc(cudaMemcpyAsync(xVector->d, x + offsetForDevice, sizeof(double) * (size_t)countForDevice, cudaMemcpyHostToDevice, st), "Initialize");
    ckc(cudaMemcpyAsync(bVector->d, b + offsetForDevice, sizeof(double) * (size_t)countForDevice, cudaMemcpyHostToDevice, st), "Initialize");
    ckc(cudaMemcpyAsync(pVector->d + offsetForDevice, xVector->d, sizeof(double) * (size_t)countForDevice, cudaMemcpyDeviceToDevice, st), "Initialize");   // :80
    int *mm = nullptr;
    ckc(cudaMalloc((void **)&mm, 2 * sizeof(int)), "Initialize");
    int init[2] = {0x7fffffff, -0x7fffffff - 1};
    ckc(cudaMemcpyAsync(mm, init, sizeof(init), cudaMemcpyHostToDevice, st), "Initialize");
    if (elementCountForDevice > 0) {
        int g = div_up(elementCountForDevice, 256); if (g > 1184) g = 1184;
        k_minmax_int<<<g, 256, 0, st>>>(elementCountForDevice, columnIndecesVector->d, mm);   // Mgcg.cu:83-84
    }
    int out[2];
    ckc(cudaMemcpyAsync(out, mm, sizeof(out), cudaMemcpyDeviceToHost, st), "Initialize");
    ckc(cudaStreamSynchronize(st), "Initialize");
    cudaFree(mm);
    *minJ = out[0]; *maxJ = out[1];
}

// ---- Mgcg.cu:88-113
void P2Host(MgcgVector *pVector, double p[], int thisCount, int thisOffset, int lastCount, int nextCount)
{
    mgcg_ctx *c = current_ctx();
    cudaStream_t st = c->stream;
    if (lastCount > 0) ckc(cudaMemcpyAsync(p + thisOffset, pVector->d + thisOffset, sizeof(double) * (size_t)lastCount, cudaMemcpyDeviceToHost, st), "P2Host");
    if (nextCount > 0) ckc(cudaMemcpyAsync(p + thisOffset + thisCount - nextCount, pVector->d + thisOffset + thisCount - nextCount, sizeof(double) * (size_t)nextCount, cudaMemcpyDeviceToHost, st), "P2Host");
    ckc(cudaStreamSynchronize(st), "P2Host");
}
void P2Device(MgcgVector *pVector, double p[], int thisCount, int thisOffset, int lastCount, int nextCount)
{
    mgcg_ctx *c = current_ctx();
    cudaStream_t st = c->stream;
    if (lastCount > 0) ckc(cudaMemcpyAsync(pVector->d + thisOffset - lastCount, p + thisOffset - lastCount, sizeof(double) * (size_t)lastCount, cudaMemcpyHostToDevice, st), "P2Device");
    if (nextCount > 0) ckc(cudaMemcpyAsync(pVector->d + thisOffset + thisCount, p + thisOffset + thisCount, sizeof(double) * (size_t)nextCount, cudaMemcpyHostToDevice, st), "P2Device");
    ckc(cudaStreamSynchronize(st), "P2Device");
}

// ---- Mgcg.cu:116-198
double Solve0(MgcgBlas *cublas, MgcgSparse *, MgcgMatDescr *, MgcgVector *elementsVector, MgcgVectorInt *rowOffsetsVector,
              MgcgVectorInt *columnIndecesVector, MgcgVector *xVector, MgcgVector *bVector, MgcgVector *ApVector,
              MgcgVector *pVector, MgcgVector *rVector, int count, int countForDevice, int offsetForDevice, int elementsCountForDevice)
{
    (void)xVector;
    mgcg_ctx *c = cublas ? cublas->ctx : current_ctx();
    // Ap = A_block * p (p global-indexed, halo filled by SyncP);  r = b - Ap;  p_block = r;  r.r
    ck(mgcg_spmv(c, countForDevice, count, elementsCountForDevice, elementsVector->d, rowOffsetsVector->d, columnIndecesVector->d,
                 pVector->d, ApVector->d, 1.0, 0.0, MGCG_SPMV_AUTO), "Solve0");
    ck(cg_init_residual(c, countForDevice, bVector->d, ApVector->d, rVector->d, pVector->d + offsetForDevice, FIN_STORE), "Solve0");
    return fetch(c, c->d_tmp_scalar);
}
double Solve1(MgcgBlas *cublas, MgcgSparse *, MgcgMatDescr *, MgcgVector *elementsVector, MgcgVectorInt *rowOffsetsVector,
              MgcgVectorInt *columnIndecesVector, MgcgVector *ApVector, MgcgVector *pVector,
              int count, int countForDevice, int offsetForDevice, int elementsCountForDevice)
{
    mgcg_ctx *c = cublas ? cublas->ctx : current_ctx();
    ck(mgcg_spmv_dot(c, countForDevice, count, elementsCountForDevice, elementsVector->d, rowOffsetsVector->d, columnIndecesVector->d,
                     pVector->d, ApVector->d, pVector->d + offsetForDevice, c->d_tmp_scalar, MGCG_SPMV_AUTO), "Solve1");
    return fetch(c, c->d_tmp_scalar);
}
double Solve2(MgcgBlas *cublas, double alpha, MgcgVector *xVector, MgcgVector *ApVector, MgcgVector *pVector, MgcgVector *rVector,
              int countForDevice, int offsetForDevice)
{
    mgcg_ctx *c = cublas ? cublas->ctx : current_ctx();
    double rr = 0.0;
    ck(mgcg_update_xr(c, countForDevice, alpha, pVector->d + offsetForDevice, ApVector->d, xVector->d, rVector->d, nullptr, &rr), "Solve2");
    return rr;
}
void Solve3(MgcgBlas *cublas, double beta, MgcgVector *pVector, MgcgVector *rVector, int countForDevice, int offsetForDevice)
{
    mgcg_ctx *c = cublas ? cublas->ctx : current_ctx();
    ck(mgcg_update_p(c, countForDevice, beta, rVector->d, pVector->d + offsetForDevice), "Solve3");
}

// ---- Mgcg.cu:201-270 / SimpleConjugateGradient.cu:12-126
static void solve_common(mgcg_ctx *c, const double *elements, const int *rowOffsets, const int *columnIndeces, double *x, const double *b,
                         double *Ap, double *p, double *r, int elementsCount, int count, double tol, int minIt, int maxIt, int strict,
                         int *iteration, double *residual)
{
    mgcg_options o;
    mgcg_options_default(&o);
    o.tolerance = tol; o.min_iteration = minIt; o.norm = MGCG_NORM_ABS_L2; o.min_strict = strict;
    // The reference's native loops never test maxIteration (Mgcg.cu:234,252; SimpleConjugateGradient.cu:86,107): only the C#
    // caller throws.  Here it is a hard stop that reports through *iteration (= max+2 iterations performed).
    o.max_iteration = maxIt;
    o.precond = g_precond; o.nx = g_nx; o.ny = g_ny; o.nz = g_nz;
    ck(mgcg_setup(c, count, elementsCount, elements, rowOffsets, columnIndeces, &o), "Solve/setup");
    double *sr = c->r, *sp = c->p, *sAp = c->Ap;
    if (r && p && Ap) { c->r = r; c->p = p; c->Ap = Ap; }   // caller-provided workspaces, like the reference
    mgcg_result res;
    int s = mgcg_solve(c, b, x, &res);
    c->r = sr; c->p = sp; c->Ap = sAp;
    if (c->graph_exec) { cudaGraphExecDestroy(c->graph_exec); c->graph_exec = nullptr; }
    if (c->graph) { cudaGraphDestroy(c->graph); c->graph = nullptr; }
    if (s != MGCG_OK && s != MGCG_ERR_NOT_CONVERGED) die("Solve");
    *iteration = res.iterations_done;
    *residual = res.residual;
}

void Solve(MgcgBlas *cublas, MgcgSparse *, MgcgMatDescr *, MgcgVector *elementsVector, MgcgVectorInt *rowOffsetsVector,
           MgcgVectorInt *columnIndecesVector, MgcgVector *xVector, MgcgVector *bVector, MgcgVector *ApVector, MgcgVector *pVector,
           MgcgVector *rVector, int elementsCount, int count, double allowableResidual, int minIteration, int maxIteration,
           int *iteration, double *residual)
{
    mgcg_ctx *c = cublas ? cublas->ctx : current_ctx();
    solve_common(c, elementsVector->d, rowOffsetsVector->d, columnIndecesVector->d, xVector->d, bVector->d,
                 ApVector->d, pVector->d, rVector->d, elementsCount, count, allowableResidual, minIteration, maxIteration, 0,
                 iteration, residual);
}

void SolveGpu_c(const double *elements, const int *rowOffsets, const int *columnIndeces, double *x, const double *b,
                int elementsCount, int count, double allowableResidual, int minIteration, int maxIteration, int *iteration, double *residual)
{
    mgcg_ctx *c = current_ctx();
    ckc(cudaMemsetAsync(x, 0, sizeof(double) * (size_t)count, c->stream), "SolveGpu");   // SimpleConjugateGradient.cu:53
    solve_common(c, elements, rowOffsets, columnIndeces, x, b, nullptr, nullptr, nullptr, elementsCount, count,
                 allowableResidual, minIteration, maxIteration, 1, iteration, residual);
}

void MgcgCompatSetPreconditioner(int precond, int nx, int ny, int nz) { g_precond = precond; g_nx = nx; g_ny = ny; g_nz = nz; }

}  // extern "C"
