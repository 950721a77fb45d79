// kernels_spmv.cu — fp64 CSR SpMV family for sm_100a with fused epilogues (p.Ap dot, residual, Jacobi sweep).
// Replaces cusparseDcsrmv (Mgcg.cu:17, SimpleConjugateGradient.cu:75,96) and, for the p.Ap epilogue, the
// host-synchronous cublasDdot that follows it (Mgcg.cu:245).  HBM-bound: 12 B per non-zero + 20 B per row.
//
//  * STREAM  (short rows, stencils): a tile of consecutive rows is a CONTIGUOUS slice of val/col, so the block
//    streams that slice with perfectly coalesced loads, stages the products a_k*x[col_k] in shared memory, and one
//    thread per row then adds its segment in STORAGE order — the same order as SparseMatrix.Multiply
//    (SparseMatrix.cs:78-85), products rounded before the add (no FMA across the staging) => bit-identical row sums.
//  * VECTOR  (long rows, e.g. the reference's 160/row band): L lanes per row, coalesced strided loads, shuffle tree.
// Both run as persistent grids (static tile->block map, so the fused reduction is deterministic).
#include "common.cuh"
#include "reduce.cuh"

namespace mgcg {

constexpr int kSpmvThreads = 256;
constexpr int kChunk = 2048;   // products staged per pass: 16 KB of shared memory

template <int EPI>
__device__ __forceinline__ void epilogue(int row, double s, const SpmvArgs &a, double &dsum)
{
    if (EPI == EPI_PLAIN) a.y[row] = s;
    else if (EPI == EPI_AXPBY) a.y[row] = (a.beta == 0.0) ? a.alpha * s : a.alpha * s + a.beta * a.y[row];
    else if (EPI == EPI_DOT) { a.y[row] = s; dsum += a.w[row] * s; }
    else if (EPI == EPI_RESID) a.y[row] = a.b[row] - s;
    else if (EPI == EPI_JACOBI || EPI == EPI_JACOBI_DOT) {
        const double bi = a.b[row];
        const double yi = (a.xrow ? a.xrow : a.x)[row] + a.omega * a.dinv[row] * (bi - s);
        a.y[row] = yi;
        if (EPI == EPI_JACOBI_DOT) dsum += bi * yi;
    }
}

template <int EPI> struct EpiTraits { static constexpr bool reduces = (EPI == EPI_DOT || EPI == EPI_JACOBI_DOT); };

// ---------------------------------------------------------------------------------------------------------
template <int EPI>
__global__ void __launch_bounds__(kSpmvThreads) k_spmv_stream(SpmvArgs a, int rows_per_tile, int num_tiles,
                                                              ReduceWs ws, Scalars *sc)
{
    if (a.gated && sc->done) return;
    __shared__ double prod[kChunk];
    __shared__ int rp[kSpmvThreads + 1];
    const int t = threadIdx.x;
    const double *__restrict__ val = a.A.val;
    const int *__restrict__ col = a.A.col;
    const double *__restrict__ x = a.x;
    double dsum = 0.0;

    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int r0 = tile * rows_per_tile;
        const int nr = min(rows_per_tile, a.A.rows - r0);
        __syncthreads();                       // previous tile's readers of rp/prod are done
        if (t < nr) rp[t] = a.A.rowptr[r0 + t];
        if (t == 0) rp[nr] = a.A.rowptr[r0 + nr];   // nr can equal blockDim.x: the end pointer needs its own load
        __syncthreads();
        const int k0 = rp[0], k1 = rp[nr];
        const int ms = (t < nr) ? rp[t] : 0, me = (t < nr) ? rp[t + 1] : 0;
        double acc = 0.0;
        for (int c0 = k0; c0 < k1; c0 += kChunk) {
            const int c1 = min(c0 + kChunk, k1);
            if (c0 != k0) __syncthreads();     // readers of the previous chunk are done
#pragma unroll 8
            for (int k = c0 + t; k < c1; k += kSpmvThreads)
                prod[k - c0] = __ldcs(val + k) * x[__ldcs(col + k)];
            __syncthreads();
            const int s0 = max(ms, c0), s1 = min(me, c1);
            for (int k = s0; k < s1; ++k) acc += prod[k - c0];
        }
        if (t < nr) epilogue<EPI>(r0 + t, acc, a, dsum);
    }
    if (EpiTraits<EPI>::reduces) {
        double m = 0.0;
        block_reduce<false>(dsum, m);
        grid_reduce_finish<false>(dsum, m, ws, a.fin, sc, a.out);
    }
}

// ---------------------------------------------------------------------------------------------------------
template <int EPI, int L>
__global__ void __launch_bounds__(kSpmvThreads) k_spmv_vector(SpmvArgs a, ReduceWs ws, Scalars *sc)
{
    if (a.gated && sc->done) return;
    const int lane = threadIdx.x & (L - 1);
    const int groups_per_block = kSpmvThreads / L;
    const int g = blockIdx.x * groups_per_block + threadIdx.x / L;
    const int ngroups = gridDim.x * groups_per_block;
    const double *__restrict__ val = a.A.val;
    const int *__restrict__ col = a.A.col;
    const double *__restrict__ x = a.x;
    const int rows = a.A.rows;
    const int rows_padded = ((rows + ngroups - 1) / ngroups) * ngroups;   // uniform trip count for the shuffles
    double dsum = 0.0;
    for (int row = g; row < rows_padded; row += ngroups) {
        double s = 0.0;
        if (row < rows) {
            const int ks = a.A.rowptr[row], ke = a.A.rowptr[row + 1];
#pragma unroll 4
            for (int k = ks + lane; k < ke; k += L) s += __ldcs(val + k) * x[__ldcs(col + k)];
        }
#pragma unroll
        for (int o = L / 2; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o, L);
        if (lane == 0 && row < rows) epilogue<EPI>(row, s, a, dsum);
    }
    if (EpiTraits<EPI>::reduces) {
        double m = 0.0;
        block_reduce<false>(dsum, m);
        grid_reduce_finish<false>(dsum, m, ws, a.fin, sc, a.out);
    }
}

// ---------------------------------------------------------------------------------------------------------
SpmvPlan spmv_make_plan(const mgcg_ctx *c, int rows, int nnz, int kernel)
{
    SpmvPlan p;
    const double avg = rows > 0 ? (double)nnz / rows : 1.0;
    int cap = c->sm_count * 8;
    if (cap > kMaxPartials) cap = kMaxPartials;
    if (kernel == MGCG_SPMV_AUTO) kernel = (avg <= 48.0) ? MGCG_SPMV_STREAM : MGCG_SPMV_VECTOR;
    p.kernel = kernel;
    if (kernel == MGCG_SPMV_STREAM) {
        int per = (int)(avg + 0.999);
        if (per < 1) per = 1;
        int r = kChunk / per;
        if (r > kSpmvThreads) r = kSpmvThreads;
        if (r < 1) r = 1;
        p.rows_per_tile = r;
        int tiles = div_up(rows > 0 ? rows : 1, r);
        p.grid = tiles < cap ? tiles : cap;
    } else {
        int L = 32;
        if (avg <= 6) L = 4; else if (avg <= 12) L = 8; else if (avg <= 24) L = 16;
        p.lanes = L;
        int need = div_up(rows > 0 ? rows : 1, kSpmvThreads / L);
        p.grid = need < cap ? need : cap;
    }
    if (p.grid < 1) p.grid = 1;
    return p;
}

template <int EPI>
static int launch_epi(mgcg_ctx *c, const SpmvPlan &plan, const SpmvArgs &a)
{
    if (a.A.rows <= 0) return MGCG_OK;
    if (plan.kernel == MGCG_SPMV_STREAM) {
        int tiles = div_up(a.A.rows, plan.rows_per_tile);
        k_spmv_stream<EPI><<<plan.grid, kSpmvThreads, 0, c->stream>>>(a, plan.rows_per_tile, tiles, c->ws, c->d_scalars);
    } else {
        switch (plan.lanes) {
        case 4: k_spmv_vector<EPI, 4><<<plan.grid, kSpmvThreads, 0, c->stream>>>(a, c->ws, c->d_scalars); break;
        case 8: k_spmv_vector<EPI, 8><<<plan.grid, kSpmvThreads, 0, c->stream>>>(a, c->ws, c->d_scalars); break;
        case 16: k_spmv_vector<EPI, 16><<<plan.grid, kSpmvThreads, 0, c->stream>>>(a, c->ws, c->d_scalars); break;
        default: k_spmv_vector<EPI, 32><<<plan.grid, kSpmvThreads, 0, c->stream>>>(a, c->ws, c->d_scalars); break;
        }
    }
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

int spmv_launch(mgcg_ctx *c, const SpmvPlan &plan, int epi, const SpmvArgs &a)
{
    switch (epi) {
    case EPI_PLAIN: return launch_epi<EPI_PLAIN>(c, plan, a);
    case EPI_AXPBY: return launch_epi<EPI_AXPBY>(c, plan, a);
    case EPI_DOT: return launch_epi<EPI_DOT>(c, plan, a);
    case EPI_RESID: return launch_epi<EPI_RESID>(c, plan, a);
    case EPI_JACOBI: return launch_epi<EPI_JACOBI>(c, plan, a);
    case EPI_JACOBI_DOT: return launch_epi<EPI_JACOBI_DOT>(c, plan, a);
    }
    set_error("spmv_launch: unknown epilogue %d", epi);
    return MGCG_ERR_INVALID;
}

}  // namespace mgcg
