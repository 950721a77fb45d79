// common.cuh — shared declarations of libMgcgGpu.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>
#include "../../include/mgcg.h"

namespace mgcg {

// ---------------------------------------------------------------------------------------------------------
// error plumbing
void set_error(const char *fmt, ...);
const char *get_error();

#define MGCG_CUDA(call)                                                                                  \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) {                                                                         \
            mgcg::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));       \
            return MGCG_ERR_CUDA;                                                                        \
        }                                                                                                \
    } while (0)

#define MGCG_CHECK(call)                                                                                 \
    do {                                                                                                 \
        int mgcg_chk_status_ = (call);                                                                   \
        if (mgcg_chk_status_ != MGCG_OK) return mgcg_chk_status_;                                        \
    } while (0)

#define MGCG_LAUNCH_CHECK() MGCG_CUDA(cudaGetLastError())

// ---------------------------------------------------------------------------------------------------------
// device-resident CG scalars: alpha/beta never visit the host inside the loop (the reference syncs twice per
// iteration to read its dots, SimpleConjugateGradient.cu:98,103).
struct Scalars {
    double rr;        // current r.r
    double rr0;       // initial r.r
    double rz;        // current r.z (== rr without preconditioner)
    double pAp;
    double alpha;
    double beta;
    double residual;  // measure of the stop rule
    double rmax;      // max|r_i| (LINF mode)
    double tolerance;
    int iteration;    // 0-based index of the iteration in flight
    int min_iteration;
    int max_iteration;
    int norm;
    int min_strict;
    int done;         // set by the device when converged / failed: every later kernel of the loop is a no-op
    int status;       // MGCG_OK / MGCG_ERR_*
    int pad;
    double rr_pend;   // single-reduce CG: r.r of the residual just updated (its stop rule is applied at the iteration's one reduction point)
    double gamma_new; // single-reduce CG: r.u of the new residual
};

// what the LAST block of a reducing kernel does with the finished sum (runs in one thread)
enum FinOp {
    FIN_STORE = 0,  // out[0] = sum
    FIN_INIT,       // rr = rr0 = rz = sum  (unpreconditioned init)
    FIN_INIT_RR,    // rr = rr0 = sum       (preconditioned init: rz comes from FIN_INIT_RZ)
    FIN_INIT_RZ,    // rz = sum
    FIN_PAP,        // pAp = sum; alpha = rz / pAp; breakdown test
    FIN_RR,         // rr_new = sum; stop rule; (no precond) beta = rr_new/rr, rz = rr = rr_new
    FIN_RR_PRECOND, // rr = sum; stop rule only
    FIN_RZ,         // beta = sum / rz; rz = sum
    FIN_STORE_MAX,  // out[0] = max
    // single-reduce (Chronopoulos-Gear) CG: the three sums of an iteration are finished where the last one lands
    FIN_CGCG_RR,        // rr_pend = sum; rmax = max
    FIN_CGCG_GAMMA,     // gamma_new = sum
    FIN_CGCG_INIT,      // delta = sum: rz = gamma_new; alpha = rz / delta; beta = 0
    FIN_CGCG_DELTA,     // delta = sum: stop rule on rr_pend; beta = gamma_new / rz; alpha = gamma_new / (delta - beta gamma_new / alpha); rz = gamma_new
    FIN_CGCG_DELTA_PLAIN // as FIN_CGCG_DELTA without preconditioner (gamma_new == rr_pend)
};

constexpr int kMaxPartials = 4096;   // upper bound on the grid of any reducing kernel

// A reduction workspace: partial sums (two lanes: sum and max), ticket counter.
struct ReduceWs {
    double *partial;     // [2 * kMaxPartials]
    unsigned int *ticket;
};

// ---------------------------------------------------------------------------------------------------------
// CSR view (device pointers, not owned)
struct Csr {
    int rows = 0, cols = 0, nnz = 0;
    const double *val = nullptr;
    const int *rowptr = nullptr;
    const int *col = nullptr;
    // coded columns (spmv_code.cu), optional: col[k] == row + dict[code[k]]; kernels that cannot read them use `col`
    const unsigned char *code = nullptr;
    const int *dict = nullptr;       // 256 ints
    int ndict = 0;                   // distinct distances
    // ... grouped into RUNS of x: distances at most kRunSpan apart share one contiguous run per tile of rows (STENCIL kernel).
    // seg_min[g] = smallest distance of run g; seg_off[c] = where code c's operand sits: g * (tile rows + kRunPad) + offset in the run
    const int *seg_min = nullptr;    // kSegMax ints
    const int *seg_off = nullptr;    // 256 ints, in units of "g * 1024 + offset" (the kernel rescales by its run length)
    int nseg = 0;                    // runs (0: too many for the STENCIL kernel)
};

// device memory behind one matrix's coded columns (spmv_code.cu): kept across set-ups, grown on demand
struct CodeStore {
    unsigned char *buf = nullptr;
    void *state = nullptr;
    size_t cap = 0;
};
constexpr int kSegMax = 12;          // x runs the STENCIL kernel stages per tile (7-point: 5, 27-point: 9)
constexpr int kRunSpan = 6;          // distances of one run lie within this span
constexpr int kRunPad = 8;           // run length = tile rows + kRunPad (span + alignment slack, even)

// SpMV plan: kernel choice and launch geometry for one matrix
struct SpmvPlan {
    int kernel = MGCG_SPMV_STREAM;  // STREAM or VECTOR
    bool tma = false;               // STREAM: use the TMA-staged variant (short rows, 16-byte aligned val/col)
    bool tma_short = false;         // ... with twice the stages of half the size (very short rows)
    int tma_wide_cap = 0;           // > 0: the WIDE variant (rows of 8..32 entries): stage capacity, ring in dynamic shared memory
    int tma_wide_smem = 0;
    int rows_per_tile = 256;        // STREAM: rows handled per tile (<= 256)
    int lanes = 8;                  // VECTOR / ROWS: lanes per row
    int rows_cap = 2552;            // ROWS: non-zeros a shared-memory stage holds (multiple of 4)
    int rows_stages = 2;            // ROWS: stages in the ring (2..8)
    int rows_unroll = 4;            // ROWS: entries a lane takes per chunk (4 or 8)
    int rows_smem = 0;              // ROWS: dynamic shared memory of the launch
    int rows_carveout = 0;          // ROWS: shared-memory share (percent) asked of the unified L1 / shared array
    int stencil_rows = 0;           // > 0: the matrix is coded with few runs — STENCIL kernel on tiles of this many rows (256 or 64)
    int grid = 0;
};

// epilogues of the SpMV kernels (s = sum_k a_ik x_k)
enum SpmvEpi {
    EPI_PLAIN = 0,   // y = s
    EPI_AXPBY,       // y = alpha*s + beta*y                     (CsrMV compat, prolongation with beta=1)
    EPI_DOT,         // y = s; partial += w_i * s                (p.Ap fused)
    EPI_RESID,       // y = b_i - s
    EPI_JACOBI,      // y = x_i + omega*dinv_i*(b_i - s)         (x is the SpMV input vector)
    EPI_JACOBI_DOT,  // as EPI_JACOBI; partial += b_i * y_i      (r.z fused into the last post-smoothing sweep)
    EPI_CHEBY,       // d = alpha*d_in_i + omega*dinv_i*(b_i - s); d_out_i = d; y = x_i + d     (one Chebyshev smoothing step)
    EPI_CHEBY_DOT    // as EPI_CHEBY; partial += b_i * y_i
};

// ---- fused halo exchange of the sharded solve (device side: halo.cuh; host side: dist.cu / peer.cu)
struct PeerCtl;
// producer side: where the boundary rows of the vector a kernel writes also go
struct Mirror {
    double *lo = nullptr;                 // lower neighbour's HIGH ghost rows: receive my rows [0, halo)
    double *hi = nullptr;                 // upper neighbour's LOW ghost rows: receive my rows [own - halo, own)
    int halo = 0, own = 0;
    unsigned long long *flag_lo = nullptr, *flag_hi = nullptr;   // xflag[me] in the neighbours' arenas
    PeerCtl *ctl = nullptr;               // my control block (null: no mirroring)
};
// consumer side: whose flags to wait for before ghost rows are read
struct GhostWait {
    PeerCtl *ctl = nullptr;               // null: no wait
    const unsigned long long *flag_lo = nullptr, *flag_hi = nullptr;   // xflag[neighbour] in MY arena
    long long clocks = 0;
};

struct SpmvArgs {
    Csr A;
    const double *x = nullptr;
    int x_lo = 0, x_hi = -1;       // index range in which x may be read (x_hi < 0: [0, A.cols)); a shifted view of a shard's vector narrows it
    const double *xrow = nullptr;  // JACOBI: x indexed by ROW (nullptr: same as x; differs when x is a shifted global-index view)
    double *y = nullptr;
    const double *b = nullptr;     // RESID / JACOBI
    const double *dinv = nullptr;  // JACOBI
    const double *w = nullptr;     // DOT
    const double *d_in = nullptr;  // CHEBY: previous direction (not read when alpha == 0)
    double *d_out = nullptr;       // CHEBY: new direction (may alias d_in: row-local)
    double alpha = 1.0, beta = 0.0, omega = 1.0;
    int fin = FIN_STORE;
    double *out = nullptr;         // FIN_STORE target
    bool gated = false;            // true: no-op once Scalars.done is set (kernels inside the CG loop)
    // red-black Gauss-Seidel half sweep (JACOBI epilogues, STREAM kernels, in place: y == x): only the rows whose grid cell has
    // parity (ix + iy + iz) & 1 == color are updated; the others keep x (and still enter the fused dot).  -1: every row.
    int color = -1;
    int cnx = 0, cny = 0;          // grid extents that decode a row index into (ix, iy, iz)
    Mirror mir;                    // sharded solve: y's boundary rows also go to the neighbours' ghost rows
    GhostWait gw;                  // sharded solve: x's ghost rows are valid once the neighbours' flags say so
};

// ---------------------------------------------------------------------------------------------------------
struct MgLevel;
struct Hierarchy;
struct DistState;

}  // namespace mgcg

// The opaque context of the C ABI.
struct mgcg_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    bool child = false;            // private context of a sharded solve (shares the parent's stream and block cache)
    int sm_count = 148;
    mgcg::Scalars *d_scalars = nullptr;
    mgcg::Scalars *h_scalars = nullptr;  // pinned, 2 slots
    mgcg::ReduceWs ws{};
    unsigned int *tail_barrier = nullptr; // {count, generation} of the persistent tail kernel's grid barrier
    double *d_tmp_scalar = nullptr;      // scratch device scalar for the standalone dot / maxabs calls
    double *h_tmp_scalar = nullptr;      // pinned
    // problem
    mgcg_options opt{};
    mgcg::Csr A;
    mgcg::SpmvPlan planA;
    bool is_setup = false;
    // work vectors (owned)
    double *r = nullptr, *p = nullptr, *Ap = nullptr;
    double *s = nullptr;           // single-reduce CG: s = A p by recurrence (allocated on first use)
    double *cgcg_u = nullptr;      // single-reduce CG: the V-cycle's output buffer (u = M r), fixed across iterations
    int graph_variant = 0;         // cg_variant the captured graph was built for
    int work_n = 0;
    mgcg::Hierarchy *mg = nullptr;
    mgcg::DistState *dist = nullptr;
    float setup_ms = 0.f;
    // solve-time accounting
    int launches = 0;
    // CUDA graph of one iteration
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t graph_exec = nullptr;
    const double *graph_b = nullptr;
    double *graph_x = nullptr;
    int graph_mode = 0;                  // 1: one iteration per launch; 2: device-side while loop (conditional node)
    int graph_body_nodes = 0;            // kernels of one iteration
    // host-pointer path buffers (mgcg_solve_host)
    void *host_path = nullptr;
    void *ell_path = nullptr;        // device system owned by mgcg_solve_host_ell (formats.cu)
    void *host_path_dist = nullptr;  // device copies owned by mgcg_solve_host_dist (dist.cu)
    mgcg::CodeStore code0;           // coded columns of the fine-level operator (spmv_code.cu)
};

namespace mgcg {

// ---- kernels_vec.cu
int vec_dot(mgcg_ctx *c, int n, const double *a, const double *b, int fin, double *out);
int vec_max_abs(mgcg_ctx *c, int n, const double *a, double *out);
int vec_axpy(mgcg_ctx *c, int n, double alpha, const double *x, double *y);
int vec_scal(mgcg_ctx *c, int n, double alpha, double *x);
int vec_copy(mgcg_ctx *c, int n, const double *x, double *y);
int vec_fill(mgcg_ctx *c, int n, double v, double *x);
// r = b - Ap ; p = r ; partial r.r  -> fin
int cg_init_residual(mgcg_ctx *c, int n, const double *b, const double *Ap, double *r, double *p, int fin, double *out = nullptr);
// x += alpha p ; r -= alpha Ap ; r.r (and max|r|) -> fin.  alpha from Scalars unless use_host_alpha.
// z0/dinv != nullptr: also z0 = omega*dinv*r (first pre-smoothing sweep of the V-cycle from a zero guess, fused)
int cg_update_xr(mgcg_ctx *c, int n, const double *p, const double *Ap, double *x, double *r, int fin,
                 bool use_host_alpha, double host_alpha, double *out,
                 double *z0 = nullptr, const double *dinv = nullptr, double omega = 0.0, const Mirror *mir = nullptr);
// single-reduce CG: p = u + beta p ; s = w + beta s ; x += alpha p ; r -= alpha s ; r.r, max|r| -> FIN_CGCG_RR  [; z0 = omega dinv r]
int cgcg_update(mgcg_ctx *c, int n, const double *u, const double *w, double *p, double *s, double *x, double *r,
                double *z0 = nullptr, const double *dinv = nullptr, double omega = 0.0, double *out = nullptr, const Mirror *mir = nullptr);
// p = z + beta p   (z == r without preconditioner)
int cg_update_p(mgcg_ctx *c, int n, const double *z, double *p, bool use_host_beta, double host_beta, const Mirror *mir = nullptr);
// x = omega * dinv .* b
int mg_jacobi_first(mgcg_ctx *c, int n, double omega, const double *dinv, const double *b, double *x, const Mirror *mir = nullptr);
// y = Minv * b for the dense coarsest level (n x n row-major)
int mg_dense_apply(mgcg_ctx *c, int n, const double *Minv, const double *b, double *y);

// ---- kernels_spmv.cu
SpmvPlan spmv_make_plan(const mgcg_ctx *c, int rows, int nnz, int kernel);
int spmv_launch(mgcg_ctx *c, const SpmvPlan &plan, int epi, const SpmvArgs &a);
int spmv_kernel_id(const SpmvPlan &plan, const Csr &A);
bool spmv_plan_reads_code(const SpmvPlan &plan);   // the kernel this plan selects understands coded columns
int spmv_plan_stencil_rows(const SpmvPlan &plan, int rows, int nnz);   // tile rows of the STENCIL kernel for this matrix shape (0: none)

// ---- spmv_code.cu
int spmv_code_build(mgcg_ctx *c, CodeStore *store, Csr *A, SpmvPlan *plan);
void spmv_code_free(CodeStore *store);

// ---- mg_setup.cu / mg_cycle.cu
int mg_build(mgcg_ctx *c);
void mg_free(mgcg_ctx *c);
void formats_free(mgcg_ctx *c);
int mg_vcycle(mgcg_ctx *c, const double *r, double **z_out, bool fuse_first_done, int rz_fin);
double *mg_level0_first_x(mgcg_ctx *c);   // buffer that receives the fused first pre-smoothing sweep
double mg_level0_first_omega(mgcg_ctx *c); // its damping: mg_omega (weighted Jacobi) or 1 (level 0 smoothed by red-black Gauss-Seidel)
const double *mg_level0_dinv(mgcg_ctx *c);

// ---- dist.cu
void dist_free(mgcg_ctx *c);
void dist_host_path_free(mgcg_ctx *c);
bool dist_x_range(const mgcg_ctx *c, int *x_lo, int *x_hi);   // sharded context: global indices [own rows +- ghost rows] a product's x may be read at

// ---- cg.cu
int cg_solve(mgcg_ctx *c, const double *b, double *x, mgcg_result *res);
int effective_max_iteration(int requested, int n);   // the option's "no limit" (>= 1 << 30 or <= 0) -> 20 n + 1000
// Build `*graph` = WHILE(!Scalars.done) { body }: `enqueue` is captured into the body of a conditional node, followed by a
// one-thread kernel that feeds the loop condition from the device-side `done` flag.  MGCG_ERR_UNSUPPORTED if the driver
// has no conditional nodes (callers fall back to one iteration per launch).
int graph_build_while(mgcg_ctx *c, int (*enqueue)(mgcg_ctx *, void *), void *arg, cudaGraph_t *graph, cudaGraphExec_t *exec, int *body_nodes);

// ---- programmatic dependent launch (PDL) of the kernels inside the CG loop: a kernel launched with the attribute may be scheduled
// while its predecessor on the stream is still draining; it does everything that reads only the (never rewritten) matrix first —
// mbarrier setup, row pointers, the TMA requests for its first tiles — and blocks in pdl_wait() before it touches anything the
// predecessor wrote (vectors, the device scalars, tickets).  Every loop kernel calls pdl_wait, so completion is transitive.
// Opt-in (MGCG_PDL=1): inside the WHILE graph the programmatic edges measured slower than plain kernel nodes (abi.cu pdl_enabled).
bool pdl_enabled();
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <typename... P, typename... A>
inline cudaError_t launch_k(bool pdl, void (*kernel)(P...), int grid, int block, size_t smem, cudaStream_t st, A... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3((unsigned)block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = (pdl && pdl_enabled()) ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, P(args)...);
}
#endif

// ---- helpers
// Allocation for the hierarchy build (alloc.cu): a size-bucketed block cache — freed blocks are re-issued to the next
// request of the same rounded size on the same stream, so a repeated build makes no driver allocation.
cudaError_t pool_alloc(void **p, size_t bytes, cudaStream_t st);
void pool_free(void *p, cudaStream_t st);
void pool_release(void *p);          // block of pool_alloc -> cache; any other device pointer -> cudaFree
void pool_reserve(size_t bytes);     // size of the next slab the block cache carves from
void pool_trim_stream(int device, cudaStream_t st);
void pool_trim(int device);          // give the cached blocks of a device (-1: all) back to the driver
size_t pool_cached_bytes();
void pool_stats(int device, long long out[5]);
int ctx_ensure_work(mgcg_ctx *c, int n);
inline int div_up(int a, int b) { return (a + b - 1) / b; }

}  // namespace mgcg
