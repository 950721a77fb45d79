// mg_cycle.cu — one multigrid V(nu1,nu2)-cycle z = M r used as the CG preconditioner.
// (No counterpart in the reference; CPU twin: oracle/mg_ref.c mg_ref_vcycle.)
// M is a fixed symmetric positive-definite linear operator: weighted-Jacobi pre/post smoothing with the same
// sweep count, R = P^T, exact (dense-inverse) coarsest solve, zero initial guess on every level.
// Per level: [x = w D^-1 b]  (nu1-1 more sweeps)  r = b - A x   b_c = R r   ...   x += P x_c   nu2 sweeps.
// The buffers used by each step are fixed, so the whole cycle can be captured into a CUDA graph.
#include "mg.cuh"

namespace mgcg {

static int jacobi_sweep(mgcg_ctx *c, MgLevel &L, double omega, bool with_dot, int fin)
{
    double *other = (L.xcur == L.xa) ? L.xb : L.xa;
    SpmvArgs a;
    a.A = L.A; a.x = L.xcur; a.y = other; a.b = L.b; a.dinv = L.dinv; a.omega = omega; a.gated = true; a.fin = fin;
    MGCG_CHECK(spmv_launch(c, L.planA, with_dot ? EPI_JACOBI_DOT : EPI_JACOBI, a));
    L.xcur = other;
    return MGCG_OK;
}

// One red-black Gauss-Seidel half sweep, in place on L.xcur: rows of parity `color` get x_i += D_ii^-1 (b_i - (A x)_i), the others
// keep x.  with_dot: sum_i b_i x_i over ALL rows goes through `fin` (r.z fused into the cycle's last half sweep).
static int gs_half_sweep(mgcg_ctx *c, MgLevel &L, int color, bool with_dot, int fin)
{
    SpmvArgs a;
    a.A = L.A; a.x = L.xcur; a.y = L.xcur; a.b = L.b; a.dinv = L.dinv; a.omega = 1.0; a.gated = true; a.fin = fin;
    a.color = color; a.cnx = L.dims[0]; a.cny = L.dims[1];
    return spmv_launch(c, L.planA, with_dot ? EPI_JACOBI_DOT : EPI_JACOBI, a);
}

// One Chebyshev smoothing step = L.cheb products (CPU twin: oracle/mg_ref.c cheby_smooth).  skip_first: step 0 from a zero guess was
// already done by the caller (x = d = c2_0 D^-1 b sits in L.xcur; its direction is read from there by step 1).
// with_dot: sum_i b_i x_i of the final x goes through `fin`.
static int cheby_smooth(mgcg_ctx *c, MgLevel &L, bool skip_first, bool with_dot, int fin)
{
    const double ratio = c->opt.mg_cheby_ratio;
    for (int k = skip_first ? 1 : 0; k < L.cheb; ++k) {
        double c1, c2;
        cheby_coeffs(L.lmax, ratio, k, &c1, &c2);
        double *other = (L.xcur == L.xa) ? L.xb : L.xa;
        const bool last = with_dot && k == L.cheb - 1;
        SpmvArgs a;
        a.A = L.A; a.x = L.xcur; a.y = other; a.b = L.b; a.dinv = L.dinv; a.gated = true; a.fin = last ? fin : FIN_STORE;
        a.alpha = c1; a.omega = c2; a.d_in = (skip_first && k == 1) ? L.xcur : L.d; a.d_out = L.d;
        MGCG_CHECK(spmv_launch(c, L.planA, last ? EPI_CHEBY_DOT : EPI_CHEBY, a));
        L.xcur = other;
    }
    return MGCG_OK;
}

// z_out receives the level-0 buffer holding z.  fuse_first_done: level 0's first sweep (xa = w D^-1 r) was already
// produced by cg_update_xr.  rz_fin: FinOp applied to sum_i r_i z_i (fused into the last post-smoothing sweep).
int mg_vcycle(mgcg_ctx *c, const double *r, double **z_out, bool fuse_first_done, int rz_fin)
{
    Hierarchy *H = c->mg;
    if (!H) { set_error("mg_vcycle: no hierarchy"); return MGCG_ERR_INVALID; }
    const int nl = (int)H->L.size();
    const double omega = c->opt.mg_omega;
    const int pre = c->opt.mg_pre_sweeps, post = c->opt.mg_post_sweeps;
    H->L[0].b = const_cast<double *>(r);

    if (nl == 1) {   // the whole problem is the dense coarsest level
        MgLevel &L = H->L[0];
        MGCG_CHECK(mg_dense_apply(c, L.A.rows, H->coarse_inv, L.b, L.xb));
        if (rz_fin >= 0) MGCG_CHECK(vec_dot(c, L.A.rows, r, L.xb, rz_fin, nullptr));
        *z_out = L.xb;
        return MGCG_OK;
    }
    // the coarse tail runs as one persistent kernel (mg_tail.cu).  It may start at level 0 only for a pure application
    // (level 0's fused first sweep and fused r.z belong to the CG kernels)
    int t0 = H->tail_first;
    if (t0 == 0 && (fuse_first_done || rz_fin >= 0)) t0 = 1;
    if (t0 > nl - 2) t0 = -1;
    const int down_end = t0 >= 0 ? t0 : nl - 1;
    // ---- down
    for (int l = 0; l < down_end; ++l) {
        MgLevel &L = H->L[l];
        L.xcur = L.xa;
        if (L.cheb) {
            double c1, c2;
            cheby_coeffs(L.lmax, c->opt.mg_cheby_ratio, 0, &c1, &c2);
            if (!(l == 0 && fuse_first_done)) MGCG_CHECK(mg_jacobi_first(c, L.A.rows, c2, L.dinv, L.b, L.xa));
            MGCG_CHECK(cheby_smooth(c, L, true, false, FIN_STORE));
            for (int s = 1; s < pre; ++s) MGCG_CHECK(cheby_smooth(c, L, false, false, FIN_STORE));
        } else {
        if (!(l == 0 && fuse_first_done)) MGCG_CHECK(mg_jacobi_first(c, L.A.rows, L.rb ? 1.0 : omega, L.dinv, L.b, L.xa));
        if (L.rb) {
            // symmetric red-black Gauss-Seidel, red then black on the way down.  From a zero guess the red half sweep is
            // x_R = D^-1 b_R; x = D^-1 b on every row followed by the black half sweep gives the same x_B = D^-1 (b_B - A_BR x_R)
            MGCG_CHECK(gs_half_sweep(c, L, 1, false, FIN_STORE));
            for (int s = 1; s < pre; ++s) { MGCG_CHECK(gs_half_sweep(c, L, 0, false, FIN_STORE)); MGCG_CHECK(gs_half_sweep(c, L, 1, false, FIN_STORE)); }
        } else {
            for (int s = 1; s < pre; ++s) MGCG_CHECK(jacobi_sweep(c, L, omega, false, FIN_STORE));
        }
        }
        SpmvArgs a;
        a.A = L.A; a.x = L.xcur; a.y = L.r; a.b = L.b; a.gated = true;
        MGCG_CHECK(spmv_launch(c, L.planA, EPI_RESID, a));
        SpmvArgs q;
        q.A = L.R.view(); q.x = L.r; q.y = H->L[l + 1].b; q.gated = true;
        MGCG_CHECK(spmv_launch(c, L.planR, EPI_PLAIN, q));
    }
    if (t0 >= 0) {
        double *xt = nullptr;
        MGCG_CHECK(mg_tail_run(c, t0, &xt));
        if (t0 == 0) { *z_out = xt; return MGCG_OK; }
    } else {   // ---- coarsest
        MgLevel &L = H->L[nl - 1];
        MGCG_CHECK(mg_dense_apply(c, L.A.rows, H->coarse_inv, L.b, L.xa));
        L.xcur = L.xa;
    }
    // ---- up
    for (int l = down_end - 1; l >= 0; --l) {
        MgLevel &L = H->L[l];
        SpmvArgs a;
        a.A = L.P.view(); a.x = H->L[l + 1].xcur; a.y = L.xcur; a.alpha = 1.0; a.beta = 1.0; a.gated = true;
        MGCG_CHECK(spmv_launch(c, L.planP, EPI_AXPBY, a));
        for (int s = 0; s < post; ++s) {
            const bool last = (l == 0 && s == post - 1 && rz_fin >= 0);
            if (L.cheb) {
                MGCG_CHECK(cheby_smooth(c, L, false, last, last ? rz_fin : FIN_STORE));
            } else if (L.rb) {      // black then red on the way up: the adjoint of the pre-smoother, so M stays symmetric
                MGCG_CHECK(gs_half_sweep(c, L, 1, false, FIN_STORE));
                MGCG_CHECK(gs_half_sweep(c, L, 0, last, last ? rz_fin : FIN_STORE));
            } else {
                MGCG_CHECK(jacobi_sweep(c, L, omega, last, last ? rz_fin : FIN_STORE));
            }
        }
    }
    if (post == 0 && rz_fin >= 0) MGCG_CHECK(vec_dot(c, H->L[0].A.rows, r, H->L[0].xcur, rz_fin, nullptr));
    *z_out = H->L[0].xcur;
    return MGCG_OK;
}

}  // namespace mgcg

using namespace mgcg;

extern "C" int mgcg_mg_apply(mgcg_ctx *c, const double *d_r, double *d_z)
{
    if (!c || !c->mg) { set_error("mgcg_mg_apply: no hierarchy (setup with precond = MGCG_PRECOND_MG)"); return MGCG_ERR_INVALID; }
    MGCG_CUDA(cudaSetDevice(c->device));
    // make sure the loop gate is open
    MGCG_CUDA(cudaMemsetAsync(c->d_scalars, 0, sizeof(Scalars), c->stream));
    double *z = nullptr;
    MGCG_CHECK(mg_vcycle(c, d_r, &z, false, -1));
    MGCG_CUDA(cudaMemcpyAsync(d_z, z, sizeof(double) * (size_t)c->A.rows, cudaMemcpyDeviceToDevice, c->stream));
    MGCG_CUDA(cudaStreamSynchronize(c->stream));
    return MGCG_OK;
}
