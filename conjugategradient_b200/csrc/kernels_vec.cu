// kernels_vec.cu — CG vector kernels: fused updates + deterministic reductions (sm_100a).
// HBM-bound streaming kernels: 128-bit loads/stores, persistent grid (multiple of the SM count), grid-stride.
// Replaces the 7-launch cuBLAS sequence of the reference loop (Mgcg.cu:244-266; SimpleConjugateGradient.cu:96-121).
#include "common.cuh"
#include "reduce.cuh"
#include "halo.cuh"

namespace mgcg {

constexpr int kVecThreads = 256;

static inline int vec_grid(const mgcg_ctx *c, long long n, int per_thread)
{
    long long need = (n + (long long)kVecThreads * per_thread - 1) / ((long long)kVecThreads * per_thread);
    long long cap = (long long)c->sm_count * 8;
    if (cap > kMaxPartials) cap = kMaxPartials;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}

static inline bool aligned16(const void *p) { return (((uintptr_t)p) & 15u) == 0; }

// ---------------------------------------------------------------------------------------------------------
// dot / max-abs
template <int V>
__global__ void __launch_bounds__(kVecThreads) k_dot(int n, const double *__restrict__ a, const double *__restrict__ b,
                                                     ReduceWs ws, int fin, Scalars *sc, double *out, bool loop_gated)
{
    pdl_wait(); pdl_launch_dependents();
    if (loop_gated && sc->done) return;
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
    grid_reduce_finish<false>(s, m, ws, fin, sc, out);
}

__global__ void __launch_bounds__(kVecThreads) k_max_abs(int n, const double *__restrict__ a, ReduceWs ws, Scalars *sc, double *out)
{
    double s = 0.0, m = 0.0;
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) m = fmax(m, fabs(a[i]));
    block_reduce<true>(s, m);
    grid_reduce_finish<true>(s, m, ws, FIN_STORE_MAX, sc, out);
}

int vec_dot(mgcg_ctx *c, int n, const double *a, const double *b, int fin, double *out)
{
    int grid = vec_grid(c, n, 4);
    bool gated = (fin == FIN_PAP || fin == FIN_RR || fin == FIN_RR_PRECOND || fin == FIN_RZ);
    if (aligned16(a) && aligned16(b))
        launch_k(gated, k_dot<2>, grid, kVecThreads, 0, c->stream, n, a, b, c->ws, fin, c->d_scalars, out, gated);
    else
        launch_k(gated, k_dot<1>, grid, kVecThreads, 0, c->stream, n, a, b, c->ws, fin, c->d_scalars, out, gated);
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

int vec_max_abs(mgcg_ctx *c, int n, const double *a, double *out)
{
    int grid = vec_grid(c, n, 4);
    k_max_abs<<<grid, kVecThreads, 0, c->stream>>>(n, a, c->ws, c->d_scalars, out);
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
// simple BLAS-1 (compat exports Axpy/Scal/Copy, Mgcg.cu:22-54)
__global__ void __launch_bounds__(kVecThreads) k_axpy(int n, double alpha, const double *__restrict__ x, double *__restrict__ y)
{
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) y[i] = y[i] + alpha * x[i];
}
__global__ void __launch_bounds__(kVecThreads) k_scal(int n, double alpha, double *__restrict__ x)
{
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) x[i] = x[i] * alpha;
}
__global__ void __launch_bounds__(kVecThreads) k_fill(int n, double v, double *__restrict__ x)
{
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) x[i] = v;
}

int vec_axpy(mgcg_ctx *c, int n, double alpha, const double *x, double *y)
{
    if (n <= 0) return MGCG_OK;
    k_axpy<<<vec_grid(c, n, 2), kVecThreads, 0, c->stream>>>(n, alpha, x, y);
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}
int vec_scal(mgcg_ctx *c, int n, double alpha, double *x)
{
    if (n <= 0) return MGCG_OK;
    k_scal<<<vec_grid(c, n, 2), kVecThreads, 0, c->stream>>>(n, alpha, x);
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}
int vec_fill(mgcg_ctx *c, int n, double v, double *x)
{
    if (n <= 0) return MGCG_OK;
    k_fill<<<vec_grid(c, n, 2), kVecThreads, 0, c->stream>>>(n, v, x);
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}
int vec_copy(mgcg_ctx *c, int n, const double *x, double *y)
{
    if (n <= 0) return MGCG_OK;
    MGCG_CUDA(cudaMemcpyAsync(y, x, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, c->stream));
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
// init: r = b - Ap ; p = r ; r.r            (ConjugateGradientCpu.cs:55-57; Mgcg.cu:226-228)
__global__ void __launch_bounds__(kVecThreads) k_init_residual(int n, const double *__restrict__ b, const double *__restrict__ Ap,
                                                               double *__restrict__ r, double *__restrict__ p,
                                                               ReduceWs ws, int fin, Scalars *sc, double *out)
{
    double s = 0.0, m = 0.0;
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        double ri = b[i] - Ap[i];
        r[i] = ri;
        if (p) p[i] = ri;
        s += ri * ri;
    }
    block_reduce<false>(s, m);
    grid_reduce_finish<false>(s, m, ws, fin, sc, out);
}

int cg_init_residual(mgcg_ctx *c, int n, const double *b, const double *Ap, double *r, double *p, int fin, double *out)
{
    k_init_residual<<<vec_grid(c, n, 2), kVecThreads, 0, c->stream>>>(n, b, Ap, r, p, c->ws, fin, c->d_scalars, out ? out : c->d_tmp_scalar);
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
// x += alpha p ; r -= alpha Ap ; r.r [; max|r|] [; z0 = omega dinv r]      one pass over 4 (6) vectors
// (ConjugateGradientCpu.cs:72-74; Solve2 Mgcg.cu:181-183).  Written as x + alpha*p and r + (-alpha)*Ap like
// LongVector.cs:48; the compiler may contract to FMA (tolerance, not bit, parity for floating point).
template <int V, bool Z0>
__global__ void __launch_bounds__(kVecThreads) k_update_xr(int n, const double *__restrict__ p, const double *__restrict__ Ap,
                                                           double *__restrict__ x, double *__restrict__ r,
                                                           ReduceWs ws, int fin, Scalars *sc, bool host_alpha, double alpha_h,
                                                           double *out, double *__restrict__ z0, const double *__restrict__ dinv,
                                                           double omega, Mirror mir)
{
    pdl_wait(); pdl_launch_dependents();
    if (!host_alpha && sc->done) return;
    bool wrote = false;
    const double alpha = host_alpha ? alpha_h : sc->alpha;
    const double nalpha = -alpha;
    double s = 0.0, m = 0.0;
    const int stride = gridDim.x * blockDim.x;
    if (V == 2) {
        const int n2 = n >> 1;
        const double2 *p2 = reinterpret_cast<const double2 *>(p);
        const double2 *A2 = reinterpret_cast<const double2 *>(Ap);
        double2 *x2 = reinterpret_cast<double2 *>(x);
        double2 *r2 = reinterpret_cast<double2 *>(r);
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
            double2 pv = p2[i], av = A2[i], xv = x2[i], rv = r2[i];
            xv.x = xv.x + alpha * pv.x; xv.y = xv.y + alpha * pv.y;
            rv.x = rv.x + nalpha * av.x; rv.y = rv.y + nalpha * av.y;
            x2[i] = xv; r2[i] = rv;
            s += rv.x * rv.x; s += rv.y * rv.y;
            m = fmax(m, fmax(fabs(rv.x), fabs(rv.y)));
            if (Z0) {
                double2 dv = reinterpret_cast<const double2 *>(dinv)[i];
                double2 zv; zv.x = omega * dv.x * rv.x; zv.y = omega * dv.y * rv.y;
                reinterpret_cast<double2 *>(z0)[i] = zv;
                if (mir.ctl) { wrote |= mirror_store(mir, 2 * i, zv.x); wrote |= mirror_store(mir, 2 * i + 1, zv.y); }
            }
        }
        if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
            int i = n - 1;
            double xi = x[i] + alpha * p[i], ri = r[i] + nalpha * Ap[i];
            x[i] = xi; r[i] = ri; s += ri * ri; m = fmax(m, fabs(ri));
            if (Z0) { const double zi = omega * dinv[i] * ri; z0[i] = zi; if (mir.ctl) wrote |= mirror_store(mir, i, zi); }
        }
    } else {
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
            double xi = x[i] + alpha * p[i], ri = r[i] + nalpha * Ap[i];
            x[i] = xi; r[i] = ri; s += ri * ri; m = fmax(m, fabs(ri));
            if (Z0) { const double zi = omega * dinv[i] * ri; z0[i] = zi; if (mir.ctl) wrote |= mirror_store(mir, i, zi); }
        }
    }
    block_reduce<true>(s, m);
    grid_reduce_finish<true>(s, m, ws, fin, sc, out);
    if (Z0) halo_signal(mir, wrote);
}

int cg_update_xr(mgcg_ctx *c, int n, const double *p, const double *Ap, double *x, double *r, int fin,
                 bool use_host_alpha, double host_alpha, double *out, double *z0, const double *dinv, double omega, const Mirror *mir)
{
    const Mirror mm = (mir && z0) ? *mir : Mirror();
    int grid = vec_grid(c, n, 2);
    bool al = aligned16(p) && aligned16(Ap) && aligned16(x) && aligned16(r) && (!z0 || (aligned16(z0) && aligned16(dinv)));
#define LAUNCH_(V, Z) launch_k(!use_host_alpha, k_update_xr<V, Z>, grid, kVecThreads, 0, c->stream, n, p, Ap, x, r, c->ws, fin, c->d_scalars, \
                               use_host_alpha, host_alpha, out, z0, dinv, omega, mm)
    if (al) { if (z0) LAUNCH_(2, true); else LAUNCH_(2, false); }
    else    { if (z0) LAUNCH_(1, true); else LAUNCH_(1, false); }
#undef LAUNCH_
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
// single-reduce (Chronopoulos-Gear) CG, one pass over 6 (8) vectors:
//   p = u + beta p ; s = w + beta s ; x += alpha p ; r -= alpha s ; r.r, max|r| [; z0 = omega dinv r]
// s tracks A p by recurrence, so the iteration's only product is w = A u and its three sums land at one point.
template <bool Z0>
__global__ void __launch_bounds__(kVecThreads) k_cgcg_update(int n, const double *__restrict__ u, const double *__restrict__ w,
                                                             double *__restrict__ p, double *__restrict__ s, double *__restrict__ x,
                                                             double *__restrict__ r, ReduceWs ws, Scalars *sc, double *out,
                                                             double *__restrict__ z0, const double *__restrict__ dinv, double omega, Mirror mir)
{
    pdl_wait(); pdl_launch_dependents();
    if (sc->done) return;
    bool wrote = false;
    const double alpha = sc->alpha, beta = sc->beta, nalpha = -alpha;
    double sum = 0.0, m = 0.0;
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const double pi = u[i] + beta * p[i];
        const double si = w[i] + beta * s[i];
        const double xi = x[i] + alpha * pi;
        const double ri = r[i] + nalpha * si;
        p[i] = pi; s[i] = si; x[i] = xi; r[i] = ri;
        sum += ri * ri;
        m = fmax(m, fabs(ri));
        if (Z0) { const double zi = omega * dinv[i] * ri; z0[i] = zi; if (mir.ctl) wrote |= mirror_store(mir, i, zi); }
    }
    block_reduce<true>(sum, m);
    grid_reduce_finish<true>(sum, m, ws, out ? FIN_STORE : FIN_CGCG_RR, sc, out);
    if (Z0) halo_signal(mir, wrote);
}

int cgcg_update(mgcg_ctx *c, int n, const double *u, const double *w, double *p, double *s, double *x, double *r,
                double *z0, const double *dinv, double omega, double *out, const Mirror *mir)
{
    const int grid = vec_grid(c, n, 2);
    const Mirror mm = (mir && z0) ? *mir : Mirror();
    if (z0) launch_k(true, k_cgcg_update<true>, grid, kVecThreads, 0, c->stream, n, u, w, p, s, x, r, c->ws, c->d_scalars, out, z0, dinv, omega, mm);
    else launch_k(true, k_cgcg_update<false>, grid, kVecThreads, 0, c->stream, n, u, w, p, s, x, r, c->ws, c->d_scalars, out, z0, dinv, omega, mm);
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
// p = z + beta p   (ConjugateGradientCpu.cs:94; Solve3 Mgcg.cu:197 does it as scal+axpy = 2 passes)
template <int V>
__global__ void __launch_bounds__(kVecThreads) k_update_p(int n, const double *__restrict__ z, double *__restrict__ p,
                                                          Scalars *sc, bool host_beta, double beta_h, Mirror mir)
{
    pdl_wait(); pdl_launch_dependents();
    if (!host_beta && sc->done) return;
    bool wrote = false;
    const double beta = host_beta ? beta_h : sc->beta;
    const int stride = gridDim.x * blockDim.x;
    if (V == 2) {
        const int n2 = n >> 1;
        const double2 *z2 = reinterpret_cast<const double2 *>(z);
        double2 *p2 = reinterpret_cast<double2 *>(p);
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
            double2 zv = z2[i], pv = p2[i];
            pv.x = zv.x + beta * pv.x; pv.y = zv.y + beta * pv.y;
            p2[i] = pv;
            if (mir.ctl) { wrote |= mirror_store(mir, 2 * i, pv.x); wrote |= mirror_store(mir, 2 * i + 1, pv.y); }
        }
        if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
            const double pi = z[n - 1] + beta * p[n - 1];
            p[n - 1] = pi;
            if (mir.ctl) wrote |= mirror_store(mir, n - 1, pi);
        }
    } else {
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
            const double pi = z[i] + beta * p[i];
            p[i] = pi;
            if (mir.ctl) wrote |= mirror_store(mir, i, pi);
        }
    }
    halo_signal(mir, wrote);
}

int cg_update_p(mgcg_ctx *c, int n, const double *z, double *p, bool use_host_beta, double host_beta, const Mirror *mir)
{
    int grid = vec_grid(c, n, 2);
    const Mirror mm = mir ? *mir : Mirror();
    if (aligned16(z) && aligned16(p)) launch_k(!use_host_beta, k_update_p<2>, grid, kVecThreads, 0, c->stream, n, z, p, c->d_scalars, use_host_beta, host_beta, mm);
    else launch_k(!use_host_beta, k_update_p<1>, grid, kVecThreads, 0, c->stream, n, z, p, c->d_scalars, use_host_beta, host_beta, mm);
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
// multigrid vector pieces
__global__ void __launch_bounds__(kVecThreads) k_jacobi_first(int n, double omega, const double *__restrict__ dinv,
                                                              const double *__restrict__ b, double *__restrict__ x, const Scalars *sc, Mirror mir)
{
    pdl_wait(); pdl_launch_dependents();
    if (sc->done) return;
    bool wrote = false;
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const double xi = omega * dinv[i] * b[i];
        x[i] = xi;
        if (mir.ctl) wrote |= mirror_store(mir, i, xi);
    }
    halo_signal(mir, wrote);
}

int mg_jacobi_first(mgcg_ctx *c, int n, double omega, const double *dinv, const double *b, double *x, const Mirror *mir)
{
    launch_k(true, k_jacobi_first, vec_grid(c, n, 2), kVecThreads, 0, c->stream, n, omega, dinv, b, x, (const Scalars *)c->d_scalars, mir ? *mir : Mirror());
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

// y = Minv b, dense n x n (n <= ~1024): one warp per row, fixed shuffle tree
__global__ void __launch_bounds__(256) k_dense_apply(int n, const double *__restrict__ Minv, const double *__restrict__ b,
                                                     double *__restrict__ y, const Scalars *sc)
{
    pdl_wait(); pdl_launch_dependents();
    if (sc->done) return;
    const int lane = threadIdx.x & 31;
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row >= n) return;
    const double *m = Minv + (size_t)row * n;
    double s = 0.0;
    for (int j = lane; j < n; j += 32) s += m[j] * b[j];
    s = warp_sum(s);
    if (lane == 0) y[row] = s;
}

int mg_dense_apply(mgcg_ctx *c, int n, const double *Minv, const double *b, double *y)
{
    int grid = div_up(n * 32, 256);
    launch_k(true, k_dense_apply, grid, 256, 0, c->stream, n, Minv, b, y, (const Scalars *)c->d_scalars);
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

}  // namespace mgcg
