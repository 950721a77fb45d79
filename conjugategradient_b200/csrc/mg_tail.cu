// mg_tail.cu — the coarse tail of the V-cycle as ONE persistent kernel.
// (No counterpart in the reference; the arithmetic is mg_cycle.cu's, step for step.)
//
// Below ~1e5 rows a multigrid level is a handful of microseconds of work per step, and a V(1,1) visit is five dependent
// steps: at 256^3 the levels under the second one are 16 dependent launches whose launch/drain gaps cost more than their
// memory traffic (and on the sharded path this tail is replicated on every rank, so it bounds the strong scaling).
// Here one grid of co-resident blocks (one per SM) walks all of those steps, separated by a grid barrier instead of a
// kernel boundary.  MEASURED (B200, 256^3, profiles/r01_tail_kernel_sweep.txt): inside a CUDA graph a kernel boundary costs
// about as much as this barrier (~2 us), and the steps themselves are bound by their dependent load chains, so the
// persistent form is NOT faster (31.0 ms with separate launches, 31.2-32.2 ms with the tail in one kernel).  It is
// therefore OFF by default (mg_tail_rows = 0) and kept as an option for launch-bound hosts (eager mode).  Vectors written inside the kernel are read back with ld.global.cg: L1 is not
// coherent across SMs and there is no kernel boundary to invalidate it; matrices, D^-1 and the dense inverse are read-only.
// Rows are dealt round-robin to groups of L lanes (L in {4,8,16,32} by the matrix's mean row length), fixed shuffle tree:
// deterministic for a given (matrix, grid).
#include "mg.cuh"
#include "reduce.cuh"
#include <string.h>

namespace mgcg {

constexpr int kTailThreads = 1024;

__device__ __forceinline__ void grid_barrier(unsigned int *bar, unsigned int nblocks)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned int gen = *((volatile unsigned int *)(bar + 1));
        if (atomicAdd(bar, 1u) == nblocks - 1) {
            *((volatile unsigned int *)bar) = 0u;
            __threadfence();
            atomicAdd(bar + 1, 1u);
        } else {
            while (*((volatile unsigned int *)(bar + 1)) == gen) { }
        }
        __threadfence();
    }
    __syncthreads();
}

enum TailOp { T_FIRST = 0, T_RESID, T_PLAIN, T_ADD, T_JACOBI };

// one step over the rows of M with groups of L lanes:  s = sum_k M_ik x[col_k]  then the step's epilogue
template <int L>
__device__ __forceinline__ void tail_rows(int op, const Csr &M, const double *x, double *y, const double *b, const double *dinv, double omega)
{
    const int lane = threadIdx.x & (L - 1);
    const int group = (blockIdx.x * blockDim.x + threadIdx.x) / L, ngroups = (gridDim.x * blockDim.x) / L;
    const int rows_padded = ((M.rows + ngroups - 1) / ngroups) * ngroups;
    for (int row = group; row < rows_padded; row += ngroups) {
        double s = 0.0;
        if (row < M.rows) {
            const int ks = M.rowptr[row], ke = M.rowptr[row + 1];
#pragma unroll 4
            for (int k = ks + lane; k < ke; k += L) s += M.val[k] * __ldcg(x + M.col[k]);
        }
#pragma unroll
        for (int o = L / 2; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o, L);
        if (lane == 0 && row < M.rows) {
            if (op == T_RESID) y[row] = __ldcg(b + row) - s;
            else if (op == T_PLAIN) y[row] = s;
            else if (op == T_ADD) y[row] = s + __ldcg(y + row);
            else y[row] = __ldcg(x + row) + omega * dinv[row] * (__ldcg(b + row) - s);    // T_JACOBI
        }
    }
}

__device__ __forceinline__ void tail_spmv(int op, int lanes, const Csr &M, const double *x, double *y, const double *b, const double *dinv, double omega)
{
    switch (lanes) {
    case 4: tail_rows<4>(op, M, x, y, b, dinv, omega); break;
    case 8: tail_rows<8>(op, M, x, y, b, dinv, omega); break;
    case 16: tail_rows<16>(op, M, x, y, b, dinv, omega); break;
    default: tail_rows<32>(op, M, x, y, b, dinv, omega); break;
    }
}

__global__ void __launch_bounds__(kTailThreads, 1) k_mg_tail(TailArgs a, const Scalars *sc)
{
    if (sc->done) return;
    const unsigned int nb = gridDim.x;
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    const int m = a.nl;
    // ---- down
    for (int l = 0; l < m - 1; ++l) {
        const TailLevel &L = a.L[l];
        double *xc = L.xa, *xo = L.xb;
        for (int i = tid; i < L.A.rows; i += nth) xc[i] = a.omega * L.dinv[i] * __ldcg(L.b + i);
        grid_barrier(a.barrier, nb);
        for (int s = 1; s < a.pre; ++s) {
            tail_spmv(T_JACOBI, L.lanesA, L.A, xc, xo, L.b, L.dinv, a.omega);
            grid_barrier(a.barrier, nb);
            double *t = xc; xc = xo; xo = t;
        }
        tail_spmv(T_RESID, L.lanesA, L.A, xc, L.r, L.b, nullptr, 0.0);
        grid_barrier(a.barrier, nb);
        tail_spmv(T_PLAIN, L.lanesR, L.R, L.r, a.L[l + 1].b, nullptr, nullptr, 0.0);
        grid_barrier(a.barrier, nb);
    }
    // ---- coarsest: x = Minv b, one warp per row
    {
        const TailLevel &Lc = a.L[m - 1];
        const int n = a.coarse_n, lane = threadIdx.x & 31, nw = nth >> 5;
        for (int row = tid >> 5; row < n; row += nw) {
            const double *mi = a.coarse_inv + (size_t)row * n;
            double s = 0.0;
            for (int j = lane; j < n; j += 32) s += mi[j] * __ldcg(Lc.b + j);
            s = warp_sum(s);
            if (lane == 0) Lc.xa[row] = s;
        }
        grid_barrier(a.barrier, nb);
    }
    // ---- up
    for (int l = m - 2; l >= 0; --l) {
        const TailLevel &L = a.L[l];
        const bool flipped = ((a.pre - 1) & 1) != 0;            // buffer that holds x after the pre-smoothing sweeps
        double *xc = flipped ? L.xb : L.xa, *xo = flipped ? L.xa : L.xb;
        const TailLevel &N = a.L[l + 1];
        const double *xn = (l + 1 == m - 1) ? N.xa : N.xfinal;
        tail_spmv(T_ADD, L.lanesP, L.P, xn, xc, nullptr, nullptr, 0.0);
        grid_barrier(a.barrier, nb);
        for (int s = 0; s < a.post; ++s) {
            tail_spmv(T_JACOBI, L.lanesA, L.A, xc, xo, L.b, L.dinv, a.omega);
            if (l > 0 || s + 1 < a.post) grid_barrier(a.barrier, nb);   // the kernel boundary orders the very last step
            double *t = xc; xc = xo; xo = t;
        }
    }
}

static int lanes_for(const Csr &M)
{
    const double avg = M.rows > 0 ? (double)M.nnz / M.rows : 1.0;
    return avg <= 6 ? 4 : avg <= 12 ? 8 : avg <= 24 ? 16 : 32;
}

// Fill H->tail for the levels [first, last]: called by mg_build once the hierarchy exists.
void mg_tail_plan(mgcg_ctx *c)
{
    Hierarchy *H = c->mg;
    H->tail_first = -1;
    const int nl = (int)H->L.size();
    if (nl < 2 || c->opt.mg_tail_rows <= 0) return;
    if (c->opt.mg_smoother != MGCG_SMOOTHER_JACOBI) return;      // the persistent tail kernel knows weighted Jacobi only
    const int limit = c->opt.mg_tail_rows;
    int first = nl - 1;
    while (first > 0 && H->L[first - 1].A.rows <= limit) --first;
    if (nl - first < 2 || nl - first > kTailMaxLevels) return;      // needs at least one sparse level above the dense one
    H->tail_first = first;
}

// Run levels [first, nl) of the cycle: H->L[first].b holds the right-hand side; on return *x_out points at level first's x.
int mg_tail_run(mgcg_ctx *c, int first, double **x_out)
{
    Hierarchy *H = c->mg;
    const int nl = (int)H->L.size();
    TailArgs a;
    memset(&a, 0, sizeof(a));
    a.nl = nl - first;
    a.omega = c->opt.mg_omega; a.pre = c->opt.mg_pre_sweeps; a.post = c->opt.mg_post_sweeps;
    a.coarse_inv = H->coarse_inv; a.coarse_n = H->coarse_n;
    a.barrier = c->tail_barrier;
    const int flips = (a.pre - 1) + a.post;
    for (int l = first; l < nl; ++l) {
        MgLevel &M = H->L[l];
        TailLevel &T = a.L[l - first];
        T.A = M.A; T.dinv = M.dinv; T.xa = M.xa; T.xb = M.xb; T.b = M.b; T.r = M.r;
        if (l + 1 < nl) {
            T.P = M.P.view(); T.R = M.R.view();
            T.lanesA = lanes_for(T.A); T.lanesP = lanes_for(T.P); T.lanesR = lanes_for(T.R);
            T.xfinal = (flips & 1) ? M.xb : M.xa;
            M.xcur = T.xfinal;
        } else {
            T.xfinal = M.xa; M.xcur = M.xa;
        }
    }
    // The blocks wait on one another at the hand-rolled grid barrier, so every one of them must be resident: size the grid from
    // the occupancy the device reports and launch cooperatively (the driver then refuses a grid that cannot be co-resident
    // instead of letting it hang behind another tenant of the device).
    int per_sm = 0;
    MGCG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_mg_tail, kTailThreads, 0));
    if (per_sm < 1) { set_error("mg_tail_run: the persistent tail kernel does not fit on this device"); return MGCG_ERR_UNSUPPORTED; }
    const int grid = c->sm_count;      // one block per SM (<= per_sm * sm_count by the check above)
    const Scalars *scp = c->d_scalars;
    void *args[] = {(void *)&a, (void *)&scp};
    MGCG_CUDA(cudaLaunchCooperativeKernel((const void *)k_mg_tail, dim3(grid), dim3(kTailThreads), args, 0, c->stream));
    c->launches++;
    *x_out = a.L[0].xfinal;
    return MGCG_OK;
}

}  // namespace mgcg
