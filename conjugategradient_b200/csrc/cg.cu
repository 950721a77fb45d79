// cg.cu — the (preconditioned) conjugate-gradient loop, device-resident scalars, speculative launch.
// Algorithm and ordering: ConjugateGradientCpu.cs:45-98 == Mgcg.cu:217-269 == SimpleConjugateGradient.cu:75-125.
// The reference reads two dots back to the host every iteration; here alpha/beta/stop-rule live in a device struct
// written by the last block of each reducing kernel, the host enqueues iterations ahead and polls a `done` flag
// with a lag (kernels enqueued past convergence see `done` and return immediately, so the result is exact).
#include "common.cuh"
#include <string.h>

namespace mgcg {

static int enqueue_iteration(mgcg_ctx *c, const double *b, double *x, std::vector<cudaEvent_t> *pev = nullptr)
{
    (void)b;
    const int n = c->A.rows;
    const bool pre = (c->opt.precond == MGCG_PRECOND_MG) && c->mg;
    // Ap = A p ; pAp ; alpha = rz / pAp                                  (ConjugateGradientCpu.cs:70-71)
    SpmvArgs a;
    a.A = c->A; a.x = c->p; a.y = c->Ap; a.w = c->p; a.fin = FIN_PAP; a.gated = true;
    if (pev) {   // profile=1: bracket the fine-level SpMV+dot launch with events on the solve stream
        cudaEvent_t a0, a1;
        MGCG_CUDA(cudaEventCreate(&a0)); MGCG_CUDA(cudaEventCreate(&a1));
        MGCG_CUDA(cudaEventRecord(a0, c->stream));
        MGCG_CHECK(spmv_launch(c, c->planA, EPI_DOT, a));
        MGCG_CUDA(cudaEventRecord(a1, c->stream));
        pev->push_back(a0); pev->push_back(a1);
    } else {
        MGCG_CHECK(spmv_launch(c, c->planA, EPI_DOT, a));
    }
    if (!pre) {
        // x += alpha p ; r -= alpha Ap ; rr' ; stop rule ; beta           (cs:72-80,93)
        MGCG_CHECK(cg_update_xr(c, n, c->p, c->Ap, x, c->r, FIN_RR, false, 0.0, nullptr));
        // p = r + beta p                                                  (cs:94)
        MGCG_CHECK(cg_update_p(c, n, c->r, c->p, false, 0.0));
    } else {
        // same update, fused with the V-cycle's first pre-smoothing sweep from a zero guess: z0 = omega D^-1 r
        MGCG_CHECK(cg_update_xr(c, n, c->p, c->Ap, x, c->r, FIN_RR_PRECOND, false, 0.0, nullptr,
                                mg_level0_first_x(c), mg_level0_dinv(c), mg_level0_first_omega(c)));
        double *z = nullptr;
        MGCG_CHECK(mg_vcycle(c, c->r, &z, true, FIN_RZ));   // z = M r ; r.z ; beta = rz'/rz (fused into the last sweep)
        MGCG_CHECK(cg_update_p(c, n, z, c->p, false, 0.0)); // p = z + beta p
    }
    return MGCG_OK;
}

// One iteration of the single-reduce (Chronopoulos-Gear) recurrence, MGCG_CG_SINGLE_REDUCE: u = M r (or r), w = A u (kept in Ap),
// s = A p by recurrence.  Same iterates as the classic loop in exact arithmetic; r.r, r.u and w.u are all finished in the last
// block of the product kernel (FIN_CGCG_DELTA), where the stop rule on the residual updated at the top of the iteration is applied.
static int enqueue_iteration_cgcg(mgcg_ctx *c, double *x)
{
    const int n = c->A.rows;
    const bool pre = (c->opt.precond == MGCG_PRECOND_MG) && c->mg;
    double *u = c->r;
    if (!pre) {
        // u == r is read (old r) and updated in the same pass: elementwise, so in place is safe
        MGCG_CHECK(cgcg_update(c, n, c->r, c->Ap, c->p, c->s, x, c->r));
    } else {
        MGCG_CHECK(cgcg_update(c, n, c->cgcg_u, c->Ap, c->p, c->s, x, c->r, mg_level0_first_x(c), mg_level0_dinv(c), mg_level0_first_omega(c)));
        MGCG_CHECK(mg_vcycle(c, c->r, &u, true, FIN_CGCG_GAMMA));
        if (u != c->cgcg_u) { set_error("cg_solve: the V-cycle's output buffer moved between iterations"); return MGCG_ERR_INVALID; }
    }
    SpmvArgs a;
    a.A = c->A; a.x = u; a.y = c->Ap; a.w = u; a.fin = pre ? FIN_CGCG_DELTA : FIN_CGCG_DELTA_PLAIN; a.gated = true;
    MGCG_CHECK(spmv_launch(c, c->planA, EPI_DOT, a));
    return MGCG_OK;
}

// "No limit" (the library default 1 << 30, or <= 0) is replaced by a finite bound: with the whole loop running as one
// device-side WHILE graph the host cannot cancel a solve that stagnates above the tolerance (a tolerance below what fp64
// can reach, a semi-definite system), so the device-side Iteration > MaxIteration test must be able to fire.  20 n + 1000
// is far above anything CG needs on an SPD system (exact arithmetic: n; the reference's own 65 536-row stress input: 2 n).
int effective_max_iteration(int requested, int n)
{
    if (requested > 0 && requested < (1 << 30)) return requested;
    const long long cap = 20LL * n + 1000;
    return (int)(cap < (1LL << 30) ? cap : (1LL << 30));
}

__global__ void k_loop_condition(cudaGraphConditionalHandle h, const Scalars *sc)
{
    if (threadIdx.x == 0) cudaGraphSetConditional(h, sc->done ? 0u : 1u);
}

int graph_build_while(mgcg_ctx *c, int (*enqueue)(mgcg_ctx *, void *), void *arg, cudaGraph_t *graph, cudaGraphExec_t *exec, int *body_nodes)
{
    cudaStream_t st = c->stream;
    cudaGraph_t g = nullptr;
    MGCG_CUDA(cudaGraphCreate(&g, 0));
    cudaGraphConditionalHandle h;
    cudaGraphNodeParams p = {};
    cudaGraphNode_t node;
    if (cudaGraphConditionalHandleCreate(&h, g, 1, cudaGraphCondAssignDefault) != cudaSuccess) { cudaGetLastError(); cudaGraphDestroy(g); return MGCG_ERR_UNSUPPORTED; }
    p.type = cudaGraphNodeTypeConditional;
    p.conditional.handle = h; p.conditional.type = cudaGraphCondTypeWhile; p.conditional.size = 1;
    if (cudaGraphAddNode(&node, g, nullptr, 0, &p) != cudaSuccess) { cudaGetLastError(); cudaGraphDestroy(g); return MGCG_ERR_UNSUPPORTED; }
    cudaGraph_t body = p.conditional.phGraph_out[0];
    if (cudaStreamBeginCaptureToGraph(st, body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
        cudaGetLastError(); cudaGraphDestroy(g); return MGCG_ERR_UNSUPPORTED;
    }
    int s = enqueue(c, arg);
    k_loop_condition<<<1, 32, 0, st>>>(h, c->d_scalars);
    cudaGraph_t got = nullptr;
    cudaError_t e = cudaStreamEndCapture(st, &got);
    if (s != MGCG_OK) { cudaGraphDestroy(g); return s; }
    if (e != cudaSuccess) { set_error("graph_build_while: capture failed: %s", cudaGetErrorString(e)); cudaGetLastError(); cudaGraphDestroy(g); return MGCG_ERR_CUDA; }
    size_t nodes = 0;
    cudaGraphGetNodes(body, nullptr, &nodes);
    e = cudaGraphInstantiate(exec, g, 0);
    if (e != cudaSuccess) { set_error("graph_build_while: instantiate failed: %s", cudaGetErrorString(e)); cudaGetLastError(); cudaGraphDestroy(g); return MGCG_ERR_UNSUPPORTED; }
    *graph = g;
    *body_nodes = (int)nodes;
    return MGCG_OK;
}

struct IterArgs { const double *b; double *x; };
static int enqueue_iteration_thunk(mgcg_ctx *c, void *arg)
{
    IterArgs *a = (IterArgs *)arg;
    if (c->opt.cg_variant == MGCG_CG_SINGLE_REDUCE) return enqueue_iteration_cgcg(c, a->x);
    return enqueue_iteration(c, a->b, a->x);
}

int cg_solve(mgcg_ctx *c, const double *b, double *x, mgcg_result *res)
{
    if (!c->is_setup) { set_error("mgcg_solve: mgcg_setup has not been called"); return MGCG_ERR_INVALID; }
    const int n = c->A.rows;
    const bool pre = (c->opt.precond == MGCG_PRECOND_MG) && c->mg;
    MGCG_CHECK(ctx_ensure_work(c, n));
    cudaStream_t st = c->stream;
    struct Events {   // destroyed on every exit path (the early returns of MGCG_CHECK / MGCG_CUDA included)
        cudaEvent_t e[4] = {nullptr, nullptr, nullptr, nullptr};
        ~Events() { for (auto &x : e) if (x) cudaEventDestroy(x); }
    } evs;
    MGCG_CUDA(cudaEventCreate(&evs.e[0]));
    MGCG_CUDA(cudaEventCreate(&evs.e[1]));
    MGCG_CUDA(cudaEventCreateWithFlags(&evs.e[2], cudaEventDisableTiming));
    MGCG_CUDA(cudaEventCreateWithFlags(&evs.e[3], cudaEventDisableTiming));
    cudaEvent_t &ev0 = evs.e[0], &ev1 = evs.e[1];
    cudaEvent_t *evp = &evs.e[2];
    c->launches = 0;

    // scalars
    Scalars h{};
    h.tolerance = c->opt.tolerance;
    h.min_iteration = c->opt.min_iteration;
    h.max_iteration = effective_max_iteration(c->opt.max_iteration, n);
    h.norm = c->opt.norm;
    h.min_strict = c->opt.min_strict;
    c->h_scalars[0] = h;
    MGCG_CUDA(cudaMemcpyAsync(c->d_scalars, &c->h_scalars[0], sizeof(Scalars), cudaMemcpyHostToDevice, st));
    MGCG_CUDA(cudaEventRecord(ev0, st));

    const bool cgcg = c->opt.cg_variant == MGCG_CG_SINGLE_REDUCE;
    if (c->opt.cg_variant != MGCG_CG_CLASSIC && !cgcg) { set_error("mgcg_solve: unknown cg_variant %d", c->opt.cg_variant); return MGCG_ERR_INVALID; }
    if (cgcg && c->opt.profile) { set_error("mgcg_solve: profile=1 times the classic loop's product kernel; not available with MGCG_CG_SINGLE_REDUCE"); return MGCG_ERR_UNSUPPORTED; }
    if (cgcg) {
        // r = b - A x ; u = M r ; w = A u ; gamma = r.u ; delta = w.u ; alpha = gamma/delta ; beta = 0 ; p = s = 0
        if (!c->s) MGCG_CUDA(cudaMalloc((void **)&c->s, sizeof(double) * (size_t)(c->work_n > 0 ? c->work_n : 1)));
        SpmvArgs a;
        a.A = c->A; a.x = x; a.y = c->Ap;
        MGCG_CHECK(spmv_launch(c, c->planA, EPI_PLAIN, a));
        double *u = c->r;
        if (!pre) {
            MGCG_CHECK(cg_init_residual(c, n, b, c->Ap, c->r, nullptr, FIN_INIT));          // rr = rr0 = rz = r.r
            MGCG_CHECK(vec_dot(c, n, c->r, c->r, FIN_CGCG_GAMMA, nullptr));                  // gamma_new = r.r (FIN_CGCG_INIT reads it)
        } else {
            MGCG_CHECK(cg_init_residual(c, n, b, c->Ap, c->r, nullptr, FIN_INIT_RR));
            MGCG_CHECK(mg_vcycle(c, c->r, &u, false, FIN_CGCG_GAMMA));
            c->cgcg_u = u;
        }
        SpmvArgs q;
        q.A = c->A; q.x = u; q.y = c->Ap; q.w = u; q.fin = FIN_CGCG_INIT;
        MGCG_CHECK(spmv_launch(c, c->planA, EPI_DOT, q));
        MGCG_CUDA(cudaMemsetAsync(c->p, 0, sizeof(double) * (size_t)n, st));
        MGCG_CUDA(cudaMemsetAsync(c->s, 0, sizeof(double) * (size_t)n, st));
    } else
    // init: Ap = A x ; r = b - Ap ; p = r (or z) ; rr                      (ConjugateGradientCpu.cs:54-57)
    {
        SpmvArgs a;
        a.A = c->A; a.x = x; a.y = c->Ap;
        MGCG_CHECK(spmv_launch(c, c->planA, EPI_PLAIN, a));
        if (!pre) {
            MGCG_CHECK(cg_init_residual(c, n, b, c->Ap, c->r, c->p, FIN_INIT));
        } else {
            MGCG_CHECK(cg_init_residual(c, n, b, c->Ap, c->r, nullptr, FIN_INIT_RR));
            double *z = nullptr;
            MGCG_CHECK(mg_vcycle(c, c->r, &z, false, FIN_INIT_RZ));
            MGCG_CHECK(vec_copy(c, n, z, c->p));
        }
    }

    // loop
    const int interval = c->opt.check_interval > 0 ? c->opt.check_interval : 8;
    const bool profile = c->opt.profile != 0;
    const bool use_graph = c->opt.use_graph && !profile;
    std::vector<cudaEvent_t> pev;
    float spmv_ms = 0.f;
    int spmv_n = 0;

    const int want_mode = use_graph ? (c->opt.use_graph >= 2 ? 2 : 1) : 0;
    if (use_graph && (!c->graph_exec || c->graph_b != b || c->graph_x != x || c->graph_variant != c->opt.cg_variant ||
                      (c->graph_mode != want_mode && c->graph_mode != -want_mode))) {
        if (c->graph_exec) { cudaGraphExecDestroy(c->graph_exec); c->graph_exec = nullptr; }
        if (c->graph) { cudaGraphDestroy(c->graph); c->graph = nullptr; }
        int saved = c->launches;
        c->graph_mode = 0;
        if (want_mode == 2) {   // the whole loop as one launch: WHILE(!done) { iteration } on the device
            IterArgs ia{b, x};
            int s = graph_build_while(c, enqueue_iteration_thunk, &ia, &c->graph, &c->graph_exec, &c->graph_body_nodes);
            if (s == MGCG_OK) c->graph_mode = 2;
            else if (s != MGCG_ERR_UNSUPPORTED) return s;
        }
        if (c->graph_mode == 0) {
            MGCG_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            IterArgs ia{b, x};
            int s = enqueue_iteration_thunk(c, &ia);
            cudaError_t e = cudaStreamEndCapture(st, &c->graph);
            if (s != MGCG_OK) return s;
            MGCG_CUDA(e);
            MGCG_CUDA(cudaGraphInstantiate(&c->graph_exec, c->graph, 0));
            size_t nodes = 0;
            cudaGraphGetNodes(c->graph, nullptr, &nodes);
            c->graph_body_nodes = (int)nodes;
            c->graph_mode = want_mode == 2 ? -2 : 1;   // -2: asked for the while loop, the driver cannot; do not retry every solve
        }
        c->graph_b = b; c->graph_x = x; c->graph_variant = c->opt.cg_variant;
        c->launches = saved;
    }
    const bool while_loop = use_graph && c->graph_mode == 2;
    long long enq = 0;
    int slot = 0;
    bool finished = false;
    bool have_prev = false;
    Scalars fin{};
    const long long hard_cap = (long long)h.max_iteration + 2;
    if (while_loop) {   // one launch: the device decides when to stop
        MGCG_CUDA(cudaGraphLaunch(c->graph_exec, st));
        finished = true;
    }
    while (!finished) {
        for (int k = 0; k < interval; ++k) {
            if (use_graph) {
                MGCG_CUDA(cudaGraphLaunch(c->graph_exec, st));
            } else {
                if (cgcg) MGCG_CHECK(enqueue_iteration_cgcg(c, x));
                else MGCG_CHECK(enqueue_iteration(c, b, x, profile ? &pev : nullptr));
            }
            ++enq;
        }
        MGCG_CUDA(cudaMemcpyAsync(&c->h_scalars[slot], c->d_scalars, sizeof(Scalars), cudaMemcpyDeviceToHost, st));
        MGCG_CUDA(cudaEventRecord(evp[slot], st));
        if (have_prev) {   // inspect the PREVIOUS batch while this one runs
            MGCG_CUDA(cudaEventSynchronize(evp[slot ^ 1]));
            if (c->h_scalars[slot ^ 1].done) { finished = true; }
        }
        have_prev = true;
        slot ^= 1;
        if (!finished && enq > hard_cap + 2LL * interval) {   // the device-side max-iteration test must have fired
            finished = true;
        }
    }
    MGCG_CUDA(cudaMemcpyAsync(&c->h_scalars[0], c->d_scalars, sizeof(Scalars), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaEventRecord(ev1, st));
    MGCG_CUDA(cudaEventSynchronize(ev1));
    fin = c->h_scalars[0];
    if (while_loop) c->launches += c->graph_body_nodes * (fin.iteration + 1);
    else if (use_graph) c->launches += (int)(c->graph_body_nodes * enq);

    float ms = 0.f;
    MGCG_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));
    for (size_t i = 0; i + 1 < pev.size(); i += 2) {
        float t = 0.f;
        cudaEventElapsedTime(&t, pev[i], pev[i + 1]);
        // launches enqueued after convergence return immediately: count only real iterations
        if ((int)(i / 2) <= fin.iteration) { spmv_ms += t; spmv_n++; }
        cudaEventDestroy(pev[i]); cudaEventDestroy(pev[i + 1]);
    }

    if (res) {
        res->status = fin.done ? fin.status : MGCG_ERR_NOT_CONVERGED;
        res->iteration = fin.iteration;
        res->iterations_done = fin.iteration + 1;
        res->residual = fin.residual;
        res->rr = fin.rr;
        res->rr0 = fin.rr0;
        res->solve_ms = ms;
        res->setup_ms = c->setup_ms;
        res->levels = mgcg_mg_levels(c);
        res->kernel_launches = c->launches;
        res->spmv_ms_total = spmv_ms;
        res->spmv_launches = spmv_n;
        res->reserved[5] = spmv_kernel_id(c->planA, c->A);   // kernel family of the fine-level SpMV (bench.py names it)
    }
    if (!fin.done) { set_error("cg_solve: iteration cap reached without the device flag"); return MGCG_ERR_NOT_CONVERGED; }
    if (fin.status != MGCG_OK) {
        set_error(fin.status == MGCG_ERR_NOT_CONVERGED ? "cg_solve: Iteration > MaxIteration (reference throws here, ConjugateGradient.cs:70-74)"
                                                       : "cg_solve: breakdown (p.Ap <= 0 or NaN)");
    }
    return fin.status;
}

}  // namespace mgcg
