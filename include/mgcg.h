/*
 * mgcg.h — C ABI of the B200-native multigrid-preconditioned conjugate-gradient library (libMgcgGpu.so).
 *
 * Two groups of entry points:
 *   (A) mgcg_*  : the native API (status codes, options struct, device- and host-pointer solves,
 *                 one-process-per-GPU sharded solve).
 *   (B) the 32 export names of the reference's MgcgGpu.dll with the reference's argument lists, so the C#
 *       P/Invoke bindings (or any other FFI) of aokomoriuta/ConjugateGradient bind unchanged.
 *       See mgcg_compat.h.
 * Plain pointers and sizes only; no C++/torch/thrust types cross this boundary.
 * Citations are file:line under the reference tree (/root/reference).
 */
#ifndef MGCG_H
#define MGCG_H

#ifdef __cplusplus
extern "C" {
#endif

#define MGCG_VERSION 200

/* ---- status codes (the reference reports no errors at all: every CUDA/cuBLAS/cuSPARSE status is dropped,
 *      Runtime.cu:23,37, Mgcg.cu:17; only the C# side throws on Iteration > MaxIteration, ConjugateGradient.cs:70-74) */
enum {
    MGCG_OK = 0,
    MGCG_ERR_INVALID = 1,        /* bad argument / not set up */
    MGCG_ERR_CUDA = 2,           /* CUDA runtime failure (see mgcg_last_error) */
    MGCG_ERR_NOT_CONVERGED = 3,  /* iteration index exceeded max_iteration (== the C# ApplicationException) */
    MGCG_ERR_BREAKDOWN = 4,      /* p.Ap <= 0 or NaN: matrix not SPD / preconditioner not SPD */
    MGCG_ERR_COMM = 5,           /* NCCL failure, or a peer-memory exchange timed out; after it from mgcg_solve_dist the set-up is void: call mgcg_setup_dist again on every rank */
    MGCG_ERR_UNSUPPORTED = 6
};

/* ---- stop rule */
enum {
    MGCG_NORM_ABS_L2 = 0,    /* sqrt(r.r) < tol      — ConjugateGradientCpu.cs:77, Mgcg.cu:251-252, SimpleConjugateGradient.cu:106-107 */
    MGCG_NORM_ABS_LINF = 1,  /* max|r_i| < tol       — Mgcg/HandmadeCL/MgcgCL/ConjugateGradientSingleGpu.cs:268 */
    MGCG_NORM_REL_L2 = 2     /* r.r / r0.r0 < tol^2  — Mgcg/ViennaCL/Mgcg/ComputerCpu.cpp:82 */
};

enum { MGCG_PRECOND_NONE = 0, MGCG_PRECOND_MG = 1 };
/* smoothers: weighted Jacobi; symmetric red-black Gauss-Seidel (two-colourable grid levels, Jacobi elsewhere); Chebyshev polynomial
 * of degree mg_cheby_degree in D^-1 A on [lmax/mg_cheby_ratio, lmax], lmax = the Gershgorin bound max_i sum_j|a_ij|/|a_ii|
 * (opt-in, SURVEY 8(f)4: fewer iterations on the 27-point variable-coefficient operator, more products per cycle) */
enum { MGCG_SMOOTHER_JACOBI = 0, MGCG_SMOOTHER_RBGS = 1, MGCG_SMOOTHER_CHEBYSHEV = 2 };
/* CG recurrence: CLASSIC = the reference's (ConjugateGradientCpu.cs:45-98: two reduction points per iteration, p.Ap then r.r / r.z);
 * SINGLE_REDUCE = Chronopoulos-Gear: the iteration's three sums (r.u, w.u, r.r) are finished at ONE point, so the sharded loop
 * has one scalar exchange per iteration instead of two.  Same iterates in exact arithmetic; opt-in, outside the parity contract */
enum { MGCG_CG_CLASSIC = 0, MGCG_CG_SINGLE_REDUCE = 1 };
enum { MGCG_SPMV_AUTO = 0, MGCG_SPMV_STREAM = 1, MGCG_SPMV_VECTOR = 2, MGCG_SPMV_ROWS = 3 };
/* internal layout of the fine operator's column indices (csrc/spmv_code.cu).  AUTO (default) = one byte per entry (the rank of
 * col - row among the matrix's distinct distances) when there are at most 256 of them, the matrix has >= 32768 rows and its SpMV
 * kernel reads codes (the TMA-staged stream kernels; with at most 8 distances the x operands are staged by TMA as well); other
 * matrices silently stay on int32 columns.  OFF = the caller's int32 columns.  Same entries, same order, same operands: every row
 * sum is bit-identical either way (a fused dot follows the launch grid of the kernel that computes it) */
enum { MGCG_CODING_AUTO = 0, MGCG_CODING_OFF = 1 };
enum { MGCG_TRANSPORT_AUTO = 0, MGCG_TRANSPORT_NCCL = 1, MGCG_TRANSPORT_PEER = 2 };
/* aggregation of the multigrid hierarchy: AUTO = geometric blocks when nx,ny,nz are given, else strength-of-connection filter +
 * MIS-2 aggregates (caller-supplied CSR without a grid, e.g. particle matrices); BLOCKS = 8 consecutive rows when there is no
 * grid (cells of cell-sorted particles); MIS2 = algebraic even when a grid hint is present */
enum { MGCG_AGG_AUTO = 0, MGCG_AGG_BLOCKS = 1, MGCG_AGG_MIS2 = 2 };

typedef struct mgcg_options {
    /* CG control — same meaning as the reference's ctor arguments (ConjugateGradient.cs:41-50) */
    double tolerance;        /* AllowableResidual */
    int    min_iteration;    /* MinIteration */
    int    max_iteration;    /* MaxIteration (honoured natively; the reference's native loops ignore it, Mgcg.cu:234) */
    int    norm;             /* MGCG_NORM_* */
    int    min_strict;       /* 1: `min < iteration` (SimpleConjugateGradient.cu:107); 0: `min <= iteration` (Mgcg.cu:252) */
    /* preconditioner */
    int    precond;          /* MGCG_PRECOND_* */
    int    nx, ny, nz;       /* optional structured-grid hint (0 = none): rows are idx = x + nx*(y + ny*z) */
    int    mg_max_levels;    /* default 12 */
    int    mg_coarse_size;   /* stop coarsening at <= this many rows; dense inverse there (default 1400) */
    int    mg_pre_sweeps;    /* default 1 */
    int    mg_post_sweeps;   /* default 1 (keep == pre for an SPD preconditioner) */
    double mg_omega;         /* Jacobi damping, default 0.85 */
    int    mg_smoother;      /* MGCG_SMOOTHER_* */
    int    mg_smoothed;      /* 1: smoothed-aggregation prolongator P=(I-w D^-1 A)T; 0: plain aggregation */
    double mg_prolong_omega; /* default 2/3 */
    /* execution */
    int    spmv_kernel;      /* MGCG_SPMV_* */
    int    check_interval;   /* host polls the device-side convergence flag every this many iterations (default 8) */
    int    use_graph;        /* 0: eager launches; 1: one iteration captured into a CUDA graph, replayed check_interval times
                                between host polls; 2 (default): the whole loop as ONE launch — a conditional WHILE node
                                whose condition is the device-side convergence flag (falls back to 1 on old drivers) */
    int    profile;          /* 1: time the fine-level SpMV launches with CUDA events (disables the graph) */
    int    mg_smoothed_levels; /* with mg_smoothed: smooth the prolongator on the finest this-many levels only, plain
                                  aggregation below (0 = every level).  Smoothing widens the coarse stencils level by level. */
    int    mg_aggregate_size;        /* geometric aggregates are s x s x s grid blocks on level 0 (default 3; needs nx,ny,nz) */
    int    mg_aggregate_size_coarse; /* the same for levels >= 1 (default 2; 0 = as level 0) */
    int    dist_replicate_below;     /* sharded solve: levels with fewer global rows than this are replicated on every
                                        rank (all-gathered right-hand side, redundant coarse solve); default 1<<18: sharding a
                                        level pays for its four halo exchanges per cycle above ~2e5 rows */
    int    dist_transport;           /* sharded solve, exchanges inside the CG loop: MGCG_TRANSPORT_AUTO (peer memory when
                                        every rank can map every other, else NCCL), _NCCL, _PEER (or fail) */
    int    mg_tail_rows;             /* > 0: multigrid levels with at most this many rows (the coarse tail of the cycle) run as
                                        ONE persistent kernel with grid barriers between the steps.  Default 0 = off: measured
                                        no faster than graph-replayed launches on B200 (mg_tail.cu) */
    int    mg_aggregation;           /* MGCG_AGG_* */
    int    mg_cheby_degree;          /* MGCG_SMOOTHER_CHEBYSHEV: products per smoothing step (default 2; 1 == Jacobi with w = 2/(lmax+lmin)) */
    double mg_strength_theta;        /* MIS-2: a_ij is a strong connection when |a_ij| >= theta * min(max_k|a_ik|, max_k|a_jk|); default 0.5 */
    double mg_cheby_ratio;           /* MGCG_SMOOTHER_CHEBYSHEV: lmax / lmin of the damped window (default 6) */
    int    cg_variant;               /* MGCG_CG_* */
    int    spmv_coding;              /* MGCG_CODING_*: coded columns of the fine operator (csrc/spmv_code.cu) */
    int    reserved[2];
} mgcg_options;

typedef struct mgcg_result {
    int    status;           /* MGCG_OK / MGCG_ERR_* */
    int    iteration;        /* 0-based index of the converging iteration == C# `Iteration` (ConjugateGradientCpu.cs:60) */
    int    iterations_done;  /* iteration + 1 == what the reference's native Solve reports (Mgcg.cu:234) */
    double residual;         /* the stop rule's measure at exit: sqrt(r.r), max|r|, or sqrt(r.r/r0.r0) */
    double rr;               /* final r.r */
    double rr0;              /* initial r.r */
    float  setup_ms;         /* hierarchy / plan construction (device time) */
    float  solve_ms;         /* CG loop (device time, CUDA events on the solve stream) */
    int    levels;           /* multigrid levels (1 = none) */
    int    kernel_launches;  /* kernels of this library launched inside the solve */
    float  spmv_ms_total;    /* profile=1: summed device time of the fine-level SpMV(+dot) launches */
    int    spmv_launches;    /* profile=1: how many of them */
    double operator_complexity; /* sum_l nnz_l / nnz_0 */
    int    reserved[6];
} mgcg_result;

typedef struct mgcg_ctx mgcg_ctx;

/* ---- lifetime */
int         mgcg_version(void);
const char *mgcg_last_error(void);                 /* thread-local message of the last failing call */
void        mgcg_options_default(mgcg_options *o); /* reference-parity defaults: ABS_L2, non-strict, no preconditioner */
mgcg_ctx   *mgcg_create(int device);               /* replaces CreateBlas/CreateSparse/CreateMatDescr (Runtime.cu:20-62) */
void        mgcg_destroy(mgcg_ctx *c);
int         mgcg_set_stream(mgcg_ctx *c, void *cuda_stream); /* run on the caller's stream (cudaStream_t) */

/* ---- solver on DEVICE pointers: replaces Solve (Mgcg.cu:201-270) and SolveGpu (SimpleConjugateGradient.cu:12-126).
 * setup analyses the matrix (SpMV plan, multigrid hierarchy built on the device); the CSR arrays are NOT copied and
 * must stay valid until the next setup / destroy.  solve: x holds the initial guess on entry (Mgcg.cu:225). */
int mgcg_setup(mgcg_ctx *c, int count, int elements_count,
               const double *d_elements, const int *d_row_offsets, const int *d_column_indeces,
               const mgcg_options *opt);
int mgcg_solve(mgcg_ctx *c, const double *d_b, double *d_x, mgcg_result *res);

/* ---- solver on HOST pointers: upload + setup + solve + download in one call (the end-to-end path) */
int mgcg_solve_host(mgcg_ctx *c, int count, int elements_count,
                    const double *elements, const int *row_offsets, const int *column_indeces,
                    const double *b, double *x, const mgcg_options *opt, mgcg_result *res);

/* ---- individual kernels on device pointers (tests, benches, and building blocks of the compat exports) */
/* y = alpha*A*x + beta*y, rectangular rows x cols — replaces CsrMV (Mgcg.cu:10-19) */
int mgcg_spmv(mgcg_ctx *c, int rows, int cols, int elements_count,
              const double *d_elements, const int *d_row_offsets, const int *d_column_indeces,
              const double *d_x, double *d_y, double alpha, double beta, int kernel);
/* y = A*x and *d_dot = sum_i w_i*y_i in one launch (fused p.Ap epilogue); deterministic */
int mgcg_spmv_dot(mgcg_ctx *c, int rows, int cols, int elements_count,
                  const double *d_elements, const int *d_row_offsets, const int *d_column_indeces,
                  const double *d_x, double *d_y, const double *d_w, double *d_dot, int kernel);
/* Coded columns of the matrix the context was set up with (an internal layout behind the CSR interface, csrc/spmv_code.cu): when the
 * matrix is coded, *ndict = the number of distinct distances col - row, dict[0..256) = the sorted distances (zero-padded),
 * codes[k] = rank of entry k's distance; otherwise *ndict = 0 and nothing else is written.
 * mgcg_spmv / mgcg_spmv_dot called with the set-up matrix's own arrays run on the coded columns as the solver does. */
int mgcg_matrix_codes(mgcg_ctx *c, int *ndict, int *dict /* [256] or NULL */, unsigned char *codes /* [elements_count] or NULL */);
/* deterministic fixed-order dot / max|.| ; result to a device scalar and (if host_out != NULL) to the host */
int mgcg_dot(mgcg_ctx *c, int n, const double *d_a, const double *d_b, double *d_out, double *host_out);
int mgcg_max_abs(mgcg_ctx *c, int n, const double *d_a, double *d_out, double *host_out);
int mgcg_axpy(mgcg_ctx *c, int n, double alpha, const double *d_x, double *d_y);  /* y += alpha x */
int mgcg_scal(mgcg_ctx *c, int n, double alpha, double *d_x);
int mgcg_copy(mgcg_ctx *c, int n, const double *d_x, double *d_y);
/* The dot product of mgcg_dot for callers that must not allocate device memory: no context, no cudaMalloc / cudaFree and no
 * allocating library routine on any call, the first included.  NO COUNTERPART IN THE REFERENCE (its dot is cublasDdot on a handle
 * that owns a workspace, Mgcg.cu:28-37); added for hosts that hand all GPU memory to their own arena at start-up, for time-step
 * loops that cannot absorb the device-wide synchronisation of cudaMalloc, and for plug-ins whose embedding application forbids
 * runtime allocation.  The reduction workspace is a 33 024-byte __device__ array inside libMgcgGpu.so (csrc/noalloc.cu); every GPU
 * that loads the library holds it until the process exits.
 *   - n <= MGCG_DOT_STATIC_MAX_N, else MGCG_ERR_INVALID with a message in mgcg_last_error (the array has room for 4096 block sums).
 *   - *d_out = sum_i d_a[i]*d_b[i], written in stream order on cuda_stream (a cudaStream_t, NULL = the default stream); the call
 *     does not synchronise.  Same summation order as mgcg_dot: same bits for the same n on the same GPU.
 *   - the first call on a device looks up the array's address there (cudaGetSymbolAddress): make it before any stream capture.
 *   - ALL calls on a GPU share the one array: calls that can overlap on the GPU (two streams, two threads) race and return
 *     garbage — order them on one stream or with events.  Each call clears the counter it needs itself. */
#define MGCG_DOT_STATIC_MAX_N (1 << 28)
int mgcg_dot_static(int device, int n, const double *d_a, const double *d_b, double *d_out, void *cuda_stream);
/* fused CG updates (replace Solve2 / Solve3 bodies, Mgcg.cu:181-183,197) */
int mgcg_update_xr(mgcg_ctx *c, int n, double alpha, const double *d_p, const double *d_Ap,
                   double *d_x, double *d_r, double *d_rr, double *host_rr);
int mgcg_update_p(mgcg_ctx *c, int n, double beta, const double *d_r, double *d_p);

/* ---- multigrid introspection (tests: hierarchy parity against the CPU twin) */
int mgcg_mg_levels(mgcg_ctx *c);
int mgcg_mg_level_size(mgcg_ctx *c, int level, int *rows, int *nnz);
/* copy level `level`'s operator (CSR) to host arrays sized by mgcg_mg_level_size */
int mgcg_mg_level_get(mgcg_ctx *c, int level, double *elements, int *row_offsets, int *column_indeces);
/* copy the fine->coarse aggregate map of `level` (rows entries) */
int mgcg_mg_level_aggregates(mgcg_ctx *c, int level, int *agg);
/* test hook: the 64-bit-checked scan that sizes every level operator, on `rows` rows of `per_row` entries (no product is built);
 * MGCG_ERR_UNSUPPORTED when the total does not fit the int32 index range of a level */
int mgcg_test_scan_counts(mgcg_ctx *c, int rows, int per_row, long long *nnz_out);
/* test hook: the block cache under the hierarchy build (csrc/alloc.cu) on `device`.  out5 = {bytes cached in the free lists,
 * slabs, their total bytes, bytes below their bump pointers, bytes of holes given back inside them}; trim != 0 first drops every
 * cached block of the device (what an out-of-memory retry does).  No counterpart in the reference (it allocates per call). */
int mgcg_test_pool_stats(int device, int trim, long long *out5);
/* z = M r : one V-cycle applied to a device vector */
int mgcg_mg_apply(mgcg_ctx *c, const double *d_r, double *d_z);

/* ---- synthetic stencil matrices assembled in device memory, one z-slab [z0, z1) of an nx*ny*nz grid at a time (SURVEY.md
 *      §8(f)2: the step before the path; the reference's generators are host loops, MgcgMain.cs:51-104).  Row layout of the
 *      reference: diagonal first (MgcgMain.cs:59-60), global 0-based columns, row offsets rebased to the slab (Mgcg.cu:73).
 *      *_nnz returns the slab's non-zero count (-1: bad arguments) so the caller can size the arrays. */
long long mgcg_generate_poisson3d7_nnz(int nx, int ny, int nz, int z0, int z1);
int mgcg_generate_poisson3d7(mgcg_ctx *c, int nx, int ny, int nz, int z0, int z1,
                             double *d_elements, int *d_row_offsets, int *d_column_indeces);
long long mgcg_generate_poisson3d27_nnz(int nx, int ny, int nz, int z0, int z1);
/* 27-point variable coefficient: d_kblk = cell coefficients on block^3 cells (ceil(nx/block)*ceil(ny/block)*ceil(nz/block)),
 * aniso3 = host array of the three axis weights */
int mgcg_generate_poisson3d27(mgcg_ctx *c, int nx, int ny, int nz, int z0, int z1, int block, const double *d_kblk,
                              const double *aniso3, double *d_elements, int *d_row_offsets, int *d_column_indeces);

/* ---- particle (MPS) pressure-Poisson matrix assembled in device memory by a cell-list neighbour search (the
 * reference receives this matrix from its particle-method caller, <= 160 entries per row, MgcgMain.cs:20).  Semantics = oracle/mps_ref.c:
 * cells of edge re, particles numbered in cell order, neighbours r < re, a_ij = -(re/r - 1), a_ii = sum_j w_ij + shift_abs, rows with
 * the diagonal first and then ascending columns.  Three calls: order the particles, size the matrix, fill it.
 * Negative return values of the two long long functions are -MGCG_ERR_*. */
long long mgcg_mps_order(mgcg_ctx *c, int n, const double *d_pos /* n x 3 */, double re, int *d_order /* n */, double *d_pos_sorted /* n x 3 */);
long long mgcg_mps_count(mgcg_ctx *c, int n, const double *d_pos_sorted, double re, int max_per_row, int *d_row_offsets /* n + 1 */);
int mgcg_mps_fill(mgcg_ctx *c, int n, const double *d_pos_sorted, double re, double shift_abs, const int *d_row_offsets,
                  double *d_elements, int *d_column_indeces);

/* ---- row-block partition rule of the reference (bit-exact): ConjugateGradientParallelGpu.cs:271-277 */
int mgcg_partition_offsets(int count, int device_count, int *offsets /* [device_count+1] */);

/* ---- one-process-per-GPU sharded solve (replaces the host-threaded ConjugateGradientParallelGpu.Solve,
 *      cs:424-565): each rank owns rows [row_offset, row_offset+local_count); column indices are GLOBAL;
 *      local row offsets are rebased to 0 (Mgcg.cu:73).  Setup traffic goes over NCCL; inside the CG loop the halo
 *      exchanges, the coarse all-gather and the scalar reductions are this library's own kernels storing into the
 *      neighbours' memory over NVLink (peer transport, one node), or NCCL calls where peers cannot be mapped. */
int mgcg_comm_unique_id(void *out128);             /* rank 0: 128-byte NCCL id to broadcast out of band */
int mgcg_comm_init(mgcg_ctx *c, const void *id128, int rank, int world);
int mgcg_setup_dist(mgcg_ctx *c, int global_count, int row_offset, int local_count, int local_elements_count,
                    const double *d_elements, const int *d_row_offsets, const int *d_column_indeces,
                    const mgcg_options *opt);
int mgcg_solve_dist(mgcg_ctx *c, const double *d_b_local, double *d_x_local, mgcg_result *res);
/* the same on HOST pointers (this rank's row block in host arrays, as the reference's multi-device class passes them to
 * Initialize, ConjugateGradientParallelGpu.cs:358-379 -> Mgcg.cu:57-85): upload + setup + solve + download in one call.
 * row_offsets may be the block's slice of the global offsets (row_offsets[0] != 0): it is rebased like Mgcg.cu:73. */
int mgcg_solve_host_dist(mgcg_ctx *c, int global_count, int row_offset, int local_count, int local_elements_count,
                         const double *elements, const int *row_offsets, const int *column_indeces,
                         const double *b_local, double *x_local, const mgcg_options *opt, mgcg_result *res);
/* collective teardown of the peer-memory transport (unmap the peers' arenas, rendezvous, free the own arena): call on every rank
 * before mgcg_destroy when the ranks are separate processes.  A later mgcg_setup_dist re-opens the transport. */
int mgcg_comm_finalize(mgcg_ctx *c);
/* collective micro-benchmark of the loop's exchanges after mgcg_setup_dist: microseconds per level-0 halo exchange,
 * per coarse all-gather and per scalar reduction (back to back, `reps` each) */
int mgcg_dist_microbench(mgcg_ctx *c, int reps, double *out_us /* [3] */);

#ifdef __cplusplus
}
#endif
#endif /* MGCG_H */
