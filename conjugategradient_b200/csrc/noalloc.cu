// noalloc.cu — mgcg_dot_static: the deterministic dot product for callers that must not allocate device memory.
//
// Who needs this.  Every other entry point of the library works through an mgcg_ctx, and mgcg_create reserves the reduction
// workspace (the per-block partial sums and the ticket counter of reduce.cuh) with cudaMalloc.  Some hosts cannot let that
// happen:
//   * simulation codes that claim all of the GPU's memory for their own arena at start-up — a cudaMalloc issued later from
//     inside a library fails, or fragments what they thought they owned;
//   * time-step loops with a latency budget — cudaMalloc / cudaFree synchronise the whole device and can take milliseconds,
//     so a convergence check inside the loop must not reach them, not even on its first use;
//   * plug-in code that runs where the embedding application forbids runtime allocation outright (its own leak / budget
//     accounting would flag memory it did not hand out).
// For them the workspace lives in the library image instead: one file-scope __device__ array, 33 024 bytes, which every GPU
// that loads libMgcgGpu.so holds from module load until the process exits whether or not this entry point is ever called.
// The caller passes the two input vectors and the output scalar; nothing else is touched.
//
// The reference has no counterpart: its dot is cublasDdot on a handle made by CreateBlas (Mgcg.cu:28-37, Runtime.cu:20-27),
// which allocates a workspace of its own.
//
// Rules of use (also in include/mgcg.h):
//   * n <= MGCG_DOT_STATIC_MAX_N (2^28).  The array holds 4096 partial sums, one per block; the limit keeps each block's share
//     at or below 65 536 products.  Larger n is refused with MGCG_ERR_INVALID and a message, never truncated.
//   * The first call on each device resolves the array's device address (cudaGetSymbolAddress, which may load the module) and
//     reads the SM count.  Make that first call outside any stream capture; later calls only enqueue a 4-byte memset and one kernel.
//   * All calls on a device share the one array.  The ticket is cleared on the caller's stream before each launch (so a launch that
//     was aborted earlier cannot poison the next), the partial sums are overwritten before they are read.  Two calls that can be
//     in flight on the same GPU at the same time — different streams, different host threads, different contexts in one process —
//     race on it and return garbage: serialise them on one stream or with events.
//   * The result lands in *d_out in stream order; the call does not synchronise.
// The summation order is mgcg_dot's (same grid rule, same block tree, same fixed-order second pass), so for a given n and GPU
// the two entry points return the same bits.
#include <mutex>

#include "common.cuh"
#include "reduce.cuh"

namespace mgcg {

constexpr int kStaticDotThreads = 256;              // == kVecThreads of kernels_vec.cu
constexpr int kStaticDotMaxN = 1 << 28;             // MGCG_DOT_STATIC_MAX_N
constexpr size_t kStaticPartialBytes = sizeof(double) * kMaxPartials;   // 32 768
constexpr size_t kStaticScratchBytes = kStaticPartialBytes + 256;       // + the ticket's own 256-byte line = 33 024

// The workspace.  In host code this name is only a handle for cudaGetSymbolAddress — it is NOT a device address and must never be
// passed to a kernel as one.
static __device__ __align__(256) unsigned char g_static_dot_scratch[kStaticScratchBytes];

template <int V>
__global__ void __launch_bounds__(kStaticDotThreads) k_dot_static(int n, const double *__restrict__ a, const double *__restrict__ b,
                                                                   double *partial, unsigned int *ticket, double *out)
{
    double s = 0.0, m = 0.0;
    const int stride = gridDim.x * blockDim.x;
    if (V == 2) {
        const int n2 = n >> 1;
        const double2 *a2 = reinterpret_cast<const double2 *>(a);
        const double2 *b2 = reinterpret_cast<const double2 *>(b);
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
            double2 u = a2[i], v = b2[i];
            s += u.x * v.x;
            s += u.y * v.y;
        }
        if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) s += a[n - 1] * b[n - 1];
    } else {
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) s += a[i] * b[i];
    }
    block_reduce<false>(s, m);
    ReduceWs ws;
    ws.partial = partial;
    ws.ticket = ticket;
    grid_reduce_finish<false>(s, m, ws, FIN_STORE, nullptr, out);
}

// per-device set-up, done once: the array's address on THAT device and its SM count
constexpr int kMaxDevices = 64;
struct StaticDotDevice {
    bool ready = false;
    unsigned char *scratch = nullptr;
    int sm_count = 0;
};
static StaticDotDevice g_static_dev[kMaxDevices];
static std::mutex g_static_mu;

static int static_dot_device(int device, StaticDotDevice **out)
{
    std::lock_guard<std::mutex> lk(g_static_mu);
    StaticDotDevice &d = g_static_dev[device];
    if (!d.ready) {
        // `device` is current here (the caller set it): the address returned is the copy of the array on this GPU
        void *p = nullptr;
        MGCG_CUDA(cudaGetSymbolAddress(&p, g_static_dot_scratch));
        if (!p) { set_error("mgcg_dot_static: cudaGetSymbolAddress returned a null address on device %d", device); return MGCG_ERR_CUDA; }
        int sm = 0;
        MGCG_CUDA(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device));
        d.scratch = static_cast<unsigned char *>(p);
        d.sm_count = sm > 0 ? sm : 148;
        d.ready = true;
    }
    *out = &d;
    return MGCG_OK;
}

}  // namespace mgcg

using namespace mgcg;

extern "C" int mgcg_dot_static(int device, int n, const double *d_a, const double *d_b, double *d_out, void *cuda_stream)
{
    if (device < 0 || device >= kMaxDevices) { set_error("mgcg_dot_static: device %d outside [0, %d)", device, kMaxDevices); return MGCG_ERR_INVALID; }
    if (n < 0) { set_error("mgcg_dot_static: negative length %d", n); return MGCG_ERR_INVALID; }
    if (n > kStaticDotMaxN) {
        set_error("mgcg_dot_static: n = %d exceeds MGCG_DOT_STATIC_MAX_N = %d (the static workspace holds %d partial sums of at most "
                  "65536 products each); use mgcg_dot on a context for longer vectors", n, kStaticDotMaxN, kMaxPartials);
        return MGCG_ERR_INVALID;
    }
    if (!d_out || (n > 0 && (!d_a || !d_b))) { set_error("mgcg_dot_static: null pointer"); return MGCG_ERR_INVALID; }
    MGCG_CUDA(cudaSetDevice(device));
    StaticDotDevice *d = nullptr;
    MGCG_CHECK(static_dot_device(device, &d));
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    double *partial = reinterpret_cast<double *>(d->scratch);
    unsigned int *ticket = reinterpret_cast<unsigned int *>(d->scratch + kStaticPartialBytes);
    // the grid rule of mgcg_dot (kernels_vec.cu vec_grid, 4 products per thread, at most 8 blocks per SM, at most kMaxPartials)
    long long need = ((long long)n + (long long)kStaticDotThreads * 4 - 1) / ((long long)kStaticDotThreads * 4);
    long long cap = (long long)d->sm_count * 8;
    if (cap > kMaxPartials) cap = kMaxPartials;
    if (need < 1) need = 1;
    const int grid = (int)(need < cap ? need : cap);
    // the ticket must be zero when the first block arrives; the partial sums need no clearing (every block writes its slot before
    // the last block reads them)
    MGCG_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st));
    const bool v2 = ((((uintptr_t)d_a) | ((uintptr_t)d_b)) & 15u) == 0;
    if (v2) k_dot_static<2><<<grid, kStaticDotThreads, 0, st>>>(n, d_a, d_b, partial, ticket, d_out);
    else    k_dot_static<1><<<grid, kStaticDotThreads, 0, st>>>(n, d_a, d_b, partial, ticket, d_out);
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}
