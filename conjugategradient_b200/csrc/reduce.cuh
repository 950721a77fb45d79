// reduce.cuh — deterministic fixed-order grid reduction in ONE launch.
// Pass 1: every block reduces its values with a fixed shuffle tree and writes partial[blockIdx.x].
// Pass 2: the block that takes the last ticket re-reads ALL partials in index order (fixed strided assignment,
// fixed tree) — the result does not depend on which block happens to be last, so it is run-to-run bit-stable
// for a given (n, grid).  Replaces cublasDdot's host-synchronous result (Mgcg.cu:35) and the OpenCL log-tree
// relaunch loop (Mgcg/HandmadeCL/MgcgCL/ConjugateGradientSingleGpu.cs:350-366).
#pragma once
#include "common.cuh"

namespace mgcg {

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_down_sync(0xffffffffu, v, o));
    return v;
}

// Sum (and max) over the block; valid in thread 0. blockDim.x must be a multiple of 32, <= 1024.
template <bool WITH_MAX>
__device__ __forceinline__ void block_reduce(double &s, double &m)
{
    __shared__ double sh_s[32];
    __shared__ double sh_m[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    s = warp_sum(s);
    if (WITH_MAX) m = warp_max(m);
    __syncthreads();   // protect sh_* against a previous use
    if (lane == 0) { sh_s[wid] = s; if (WITH_MAX) sh_m[wid] = m; }
    __syncthreads();
    if (wid == 0) {
        s = lane < nw ? sh_s[lane] : 0.0;
        s = warp_sum(s);
        if (WITH_MAX) { m = lane < nw ? sh_m[lane] : 0.0; m = warp_max(m); }
    }
}

// The reference's stop rule on the residual just updated (rr_new = r.r, mx = max|r|): sets residual / rmax / rr, and either
// done (+ status) or iteration + 1.  Returns true when the loop ends here.
__device__ __forceinline__ bool stop_rule(double rr_new, double mx, Scalars *sc)
{
    double res;
    bool below;
    if (sc->norm == MGCG_NORM_ABS_L2) { res = sqrt(rr_new); below = res < sc->tolerance; }
    else if (sc->norm == MGCG_NORM_ABS_LINF) { res = mx; below = res < sc->tolerance; }
    else if (sc->rr0 == 0.0) { res = 0.0; below = true; }
    else { res = sqrt(rr_new / sc->rr0); below = (rr_new / sc->rr0) < sc->tolerance * sc->tolerance; }
    sc->residual = res;
    sc->rmax = mx;
    const int it = sc->iteration;
    bool conv;
    if (sc->min_strict) {                       // SimpleConjugateGradient.cu:107
        if (it > sc->max_iteration) { sc->status = MGCG_ERR_NOT_CONVERGED; sc->done = 1; conv = true; }
        else conv = (sc->min_iteration < it) && below;
    } else {                                    // ConjugateGradient.cs:64-77
        if (it < sc->min_iteration) conv = false;
        else if (it > sc->max_iteration) { sc->status = MGCG_ERR_NOT_CONVERGED; sc->done = 1; conv = true; }
        else conv = below;
    }
    if (!(rr_new == rr_new)) { sc->status = MGCG_ERR_BREAKDOWN; sc->done = 1; conv = true; }   // NaN
    sc->rr = rr_new;
    if (conv) sc->done = 1;
    else sc->iteration = it + 1;
    return conv;
}

// what the last block does with the finished sums
__device__ __forceinline__ void finalize(int fin, double sum, double mx, Scalars *sc, double *out)
{
    switch (fin) {
    case FIN_STORE: out[0] = sum; break;
    case FIN_STORE_MAX: out[0] = mx; break;
    case FIN_INIT: sc->rr = sum; sc->rr0 = sum; sc->rz = sum; break;
    case FIN_INIT_RR: sc->rr = sum; sc->rr0 = sum; break;
    case FIN_INIT_RZ: sc->rz = sum; break;
    case FIN_PAP: {
        sc->pAp = sum;
        sc->alpha = (sc->rz == 0.0) ? 0.0 : sc->rz / sum;   // r == 0: already solved, stay finite
        if (!(sum > 0.0) && sc->rz != 0.0) { sc->status = MGCG_ERR_BREAKDOWN; sc->done = 1; }
        break;
    }
    case FIN_RR:
    case FIN_RR_PRECOND: {
        const double rr_old = sc->rr;
        if (!stop_rule(sum, mx, sc) && fin == FIN_RR) { sc->beta = (rr_old == 0.0) ? 0.0 : sum / rr_old; sc->rz = sum; }   // ConjugateGradientCpu.cs:93
        break;
    }
    case FIN_CGCG_RR: sc->rr_pend = sum; sc->rmax = mx; break;
    case FIN_CGCG_GAMMA: sc->gamma_new = sum; break;
    case FIN_CGCG_INIT: {
        sc->rz = sc->gamma_new;
        sc->pAp = sum;
        sc->alpha = (sc->rz == 0.0) ? 0.0 : sc->rz / sum;
        sc->beta = 0.0;
        if (!(sum > 0.0) && sc->rz != 0.0) { sc->status = MGCG_ERR_BREAKDOWN; sc->done = 1; }
        break;
    }
    case FIN_CGCG_DELTA:
    case FIN_CGCG_DELTA_PLAIN: {
        if (stop_rule(sc->rr_pend, sc->rmax, sc)) break;
        const double g = (fin == FIN_CGCG_DELTA) ? sc->gamma_new : sc->rr_pend;
        const double beta = (sc->rz == 0.0) ? 0.0 : g / sc->rz;
        const double den = sum - beta * g / sc->alpha;      // == p'.A p' of the next direction (Chronopoulos & Gear 1989)
        sc->beta = beta;
        sc->pAp = den;
        sc->alpha = (g == 0.0) ? 0.0 : g / den;
        sc->rz = g;
        if (!(den > 0.0) && g != 0.0) { sc->status = MGCG_ERR_BREAKDOWN; sc->done = 1; }
        break;
    }
    case FIN_RZ: sc->beta = (sc->rz == 0.0) ? 0.0 : sum / sc->rz; sc->rz = sum; break;
    }
}

// Grid-level completion. Call from ALL threads of every block with the block's value in thread 0.
template <bool WITH_MAX>
__device__ __forceinline__ void grid_reduce_finish(double s, double m, ReduceWs ws, int fin, Scalars *sc, double *out)
{
    __shared__ bool is_last;
    if (threadIdx.x == 0) {
        ws.partial[blockIdx.x] = s;
        if (WITH_MAX) ws.partial[kMaxPartials + blockIdx.x] = m;
        __threadfence();
        unsigned int t = atomicAdd(ws.ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) {   // fixed assignment, ascending
        a += __ldcg(&ws.partial[i]);
        if (WITH_MAX) b = fmax(b, __ldcg(&ws.partial[kMaxPartials + i]));
    }
    block_reduce<WITH_MAX>(a, b);
    if (threadIdx.x == 0) {
        *ws.ticket = 0u;   // ready for the next launch on this stream
        finalize(fin, a, b, sc, out);
    }
}

}  // namespace mgcg
