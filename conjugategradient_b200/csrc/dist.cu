// dist.cu — one-process-per-GPU sharded MGCG over NCCL (see dist.cuh for the layout).
#include "dist.cuh"
#include "reduce.cuh"
#include <dlfcn.h>
#include <mutex>
#include <string.h>

namespace mgcg {

// ---------------------------------------------------------------------------------------------------------
// NCCL is bound at run time: the library must load (and export every symbol) on a machine without NCCL or a GPU,
// and a host process that already carries NCCL (e.g. torch) shares its copy through the common soname.
int nccl_load(NcclApi **out)
{
    static NcclApi api;
    static std::mutex mu;
    std::lock_guard<std::mutex> lk(mu);
    if (!api.handle) {
        void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) { set_error("cannot load libnccl.so.2: %s", dlerror()); return MGCG_ERR_COMM; }
#define SYM_(field, name)                                                                     \
    *(void **)(&api.field) = dlsym(h, name);                                                  \
    if (!api.field) { set_error("libnccl: missing symbol %s", name); dlclose(h); return MGCG_ERR_COMM; }
        SYM_(GetUniqueId, "ncclGetUniqueId") SYM_(CommInitRank, "ncclCommInitRank") SYM_(CommDestroy, "ncclCommDestroy")
        SYM_(AllReduce, "ncclAllReduce") SYM_(AllGather, "ncclAllGather") SYM_(Broadcast, "ncclBroadcast")
        SYM_(Send, "ncclSend") SYM_(Recv, "ncclRecv") SYM_(GroupStart, "ncclGroupStart") SYM_(GroupEnd, "ncclGroupEnd")
        SYM_(GetErrorString, "ncclGetErrorString")
#undef SYM_
        *(void **)(&api.CommInitRankConfig) = dlsym(h, "ncclCommInitRankConfig");
        api.handle = h;
    }
    *out = &api;
    return MGCG_OK;
}

static void dist_graph_free(DistState *D);
static int dist_level_chebyshev(mgcg_ctx *c, DistLevel &L);

// one int from device memory, ordered after the work queued on the context's (non-blocking) stream
static int fetch_int(mgcg_ctx *c, int *h, const int *d)
{
    MGCG_CUDA(cudaMemcpyAsync(h, d, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    MGCG_CUDA(cudaStreamSynchronize(c->stream));
    return MGCG_OK;
}


void dist_free_levels(DistState *D)
{
    for (size_t l = 0; l < D->L.size(); ++l) {
        DistLevel &L = D->L[l];
        L.Aown.free_all(); L.Pext.free_all(); L.R.free_all();
        if (L.dinv) pool_release(L.dinv);
        if (!D->vec_in_arena) {     // (arena vectors belong to the peer transport)
            if (L.xa) pool_release(L.xa);
            if (L.xb) pool_release(L.xb);
            if (L.r) pool_release(L.r);
        }
        if (L.d) pool_release(L.d);
        if (l > 0 && L.b) pool_release(L.b);
    }
    D->L.clear();
    if (D->coarse) {     // kept for the next setup: creating a context costs pinned allocations and a dozen driver calls
        if (D->coarse_scalars_orig) D->coarse->d_scalars = D->coarse_scalars_orig;
        D->coarse_scalars_orig = nullptr;
        mg_free(D->coarse);
        if (D->coarse_spare) mgcg_destroy(D->coarse_spare);
        D->coarse_spare = D->coarse; D->coarse = nullptr;
    }
    D->coarse_A.free_all();
    if (D->coarse_b) { pool_release(D->coarse_b); D->coarse_b = nullptr; }
    if (D->p && !D->vec_in_arena) pool_release(D->p);
    if (D->r && !D->vec_in_arena) pool_release(D->r);
    D->p = D->r = nullptr;
    D->vec_in_arena = false;
    if (D->Ap) { pool_release(D->Ap); D->Ap = nullptr; }
    if (D->s) { pool_release(D->s); D->s = nullptr; }
}

void dist_free(mgcg_ctx *c)
{
    DistState *D = c->dist;
    if (!D) return;
    dist_graph_free(D);
    peer_close(c, false);
    dist_free_levels(D);
    if (D->coarse_spare) { mgcg_destroy(D->coarse_spare); D->coarse_spare = nullptr; }
    if (D->partial_send) cudaFree(D->partial_send);
    if (D->partial_all) cudaFree(D->partial_all);
    if (D->gather_ints) cudaFree(D->gather_ints);
    if (D->comm && D->api) D->api->CommDestroy(D->comm);
    delete D;
    c->dist = nullptr;
}

// Exchange the ghost rows of a [lo|own|hi] vector with the two neighbouring ranks (replaces SyncP = P2Host + P2Device
// through a host buffer, ConjugateGradientParallelGpu.cs:384-419).  Every rank sends `halo` rows each way.
int dist_halo_exchange(mgcg_ctx *c, double *v, int own, int lo, int hi, int halo)
{
    DistState *D = c->dist;
    if (D->world == 1 || halo <= 0) return MGCG_OK;
    D->halo_exchanges++;
    if (D->peer.active) return peer_halo_exchange(c, v, own, lo, hi, halo);
    NcclApi *a = D->api;
    double *own0 = v + lo;
    MGCG_NCCL(a, GroupStart());
    if (D->rank > 0) {
        MGCG_NCCL(a, Send(own0, (size_t)halo, ncclDouble, D->rank - 1, D->comm, c->stream));
        MGCG_NCCL(a, Recv(v, (size_t)lo, ncclDouble, D->rank - 1, D->comm, c->stream));
    }
    if (D->rank < D->world - 1) {
        MGCG_NCCL(a, Send(own0 + own - halo, (size_t)halo, ncclDouble, D->rank + 1, D->comm, c->stream));
        MGCG_NCCL(a, Recv(own0 + own, (size_t)hi, ncclDouble, D->rank + 1, D->comm, c->stream));
    }
    MGCG_NCCL(a, GroupEnd());
    return MGCG_OK;
}


// ---------------------------------------------------------------------------------------------------------
// setup helpers
__global__ void k_halfband(Csr A, int row_base, int *out)
{
    int m = 0;
    const int nz0 = A.rowptr[0];
    (void)nz0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.rows; i += gridDim.x * blockDim.x)
        for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k) m = max(m, abs(A.col[k] - (i + row_base)));
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_down_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(out, m);
}

// the aggregate (global coarse row) of global fine row i: aggregates are local to the owning rank's block
__device__ __forceinline__ int agg_of(const LevelGeom &g, int i)
{
    int r = 0;
    while (r + 1 < g.world && i >= g.off[r + 1]) ++r;
    const int li = i - g.off[r];
    if (!g.structured) return g.coff[r] + li / g.agg;
    const int x = li % g.nx, y = (li / g.nx) % g.ny, z = li / (g.nx * g.ny);
    return g.coff[r] + (x / g.agg) + g.cx * ((y / g.agg) + g.cy * (z / g.agg));
}
__global__ void k_fill_agg(LevelGeom g, int g0, int count, int *agg)
{
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < count; t += gridDim.x * blockDim.x) agg[t] = agg_of(g, g0 + t);
}

// row pointers of the extended block [ghost rows from r-1 | own rows | ghost rows from r+1]
__global__ void k_ext_rowptr(int e_lo, int own, int e_hi, const int *rp_lo, const int *rp_own, const int *rp_hi,
                             int nnz_lo, int nnz_own, int *rp_ext)
{
    const int total = e_lo + own + e_hi;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= total; i += gridDim.x * blockDim.x) {
        int v;
        if (i < e_lo) v = rp_lo[i] - rp_lo[0];
        else if (i <= e_lo + own) v = rp_own[i - e_lo] - rp_own[0] + nnz_lo;
        else v = rp_hi[i - e_lo - own] - rp_hi[0] + nnz_lo + nnz_own;
        rp_ext[i] = v;
    }
}

// transpose of a row block of P restricted to the owned coarse columns [coff, coff+cown); everything else goes to a
// dump row `cown` that the caller drops
__global__ void k_transpose_expand_filtered(Csr P, int coff, int cown, unsigned long long *keys, double *vals)
{
    const int nz0 = P.rowptr[0];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P.rows; i += gridDim.x * blockDim.x)
        for (int k = P.rowptr[i]; k < P.rowptr[i + 1]; ++k) {
            const int cc = P.col[k] - coff;
            const unsigned row = (cc >= 0 && cc < cown) ? (unsigned)cc : (unsigned)cown;
            keys[k - nz0] = ((unsigned long long)row << 32) | (unsigned)i;
            vals[k - nz0] = P.val[k];
        }
}

__global__ void k_add_int(int n, int *a, int v)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) a[i] += v;
}

static inline int gridn(long long n) { long long g = (n + 255) / 256; if (g < 1) g = 1; if (g > 65536) g = 65536; return (int)g; }

// fill L.geom (aggregate rule of level L -> next level) from the ranks' row offsets of this level
static void make_geom(DistLevel &L, int world, const int *offs, int nx, int ny, int a_struct)
{
    LevelGeom &g = L.geom;
    memset(&g, 0, sizeof(g));
    g.world = world;
    for (int r = 0; r <= world; ++r) g.off[r] = offs[r];
    bool structured = nx > 0 && ny > 0;
    if (structured) for (int r = 0; r < world; ++r) if ((offs[r + 1] - offs[r]) % (nx * ny) != 0) structured = false;
    g.structured = structured ? 1 : 0;
    g.coff[0] = 0;
    if (structured) {
        g.nx = nx; g.ny = ny; g.agg = a_struct;
        g.cx = (nx + a_struct - 1) / a_struct; g.cy = (ny + a_struct - 1) / a_struct;
        for (int r = 0; r < world; ++r) {
            const int nzl = (offs[r + 1] - offs[r]) / (nx * ny);
            g.coff[r + 1] = g.coff[r] + g.cx * g.cy * ((nzl + a_struct - 1) / a_struct);
        }
    } else {
        g.agg = 8;
        for (int r = 0; r < world; ++r) g.coff[r + 1] = g.coff[r] + (offs[r + 1] - offs[r] + 7) / 8;
    }
}

static int level_halfband(mgcg_ctx *c, const Csr &A, int row_base, int *out)
{
    DistState *D = c->dist;
    int *d = nullptr;
    MGCG_CUDA(pool_alloc((void **)&d, sizeof(int), c->stream));
    MGCG_CUDA(cudaMemsetAsync(d, 0, sizeof(int), c->stream));
    if (A.rows > 0) { k_halfband<<<gridn(A.rows), 256, 0, c->stream>>>(A, row_base, d); MGCG_LAUNCH_CHECK(); }
    // max over ranks: all-gather + host max (the setup uses two NCCL connection classes only, the collective ring and
    // point-to-point, both opened once in mgcg_comm_init; an all-reduce would open its own on first use)
    MGCG_NCCL(D->api, AllGather(d, D->gather_ints, 1, ncclInt, D->comm, c->stream));
    int all[kMaxRanks];
    MGCG_CUDA(cudaMemcpyAsync(all, D->gather_ints, sizeof(int) * (size_t)D->world, cudaMemcpyDeviceToHost, c->stream));
    MGCG_CUDA(cudaStreamSynchronize(c->stream));
    pool_release(d);
    int m = 0;
    for (int r = 0; r < D->world; ++r) m = all[r] > m ? all[r] : m;
    *out = m;
    return MGCG_OK;
}

static int alloc_level_vectors_dist(DistLevel &L, bool level0, cudaStream_t st)
{
    const size_t e = (size_t)L.lo + L.own + L.hi + 1;
    MGCG_CUDA(pool_alloc((void **)&L.xa, sizeof(double) * e, st));
    MGCG_CUDA(pool_alloc((void **)&L.xb, sizeof(double) * e, st));
    MGCG_CUDA(cudaMemsetAsync(L.xa, 0, sizeof(double) * e, st));
    MGCG_CUDA(cudaMemsetAsync(L.xb, 0, sizeof(double) * e, st));
    MGCG_CUDA(pool_alloc((void **)&L.r, sizeof(double) * e, st));
    MGCG_CUDA(cudaMemsetAsync(L.r, 0, sizeof(double) * e, st));
    if (!level0) MGCG_CUDA(pool_alloc((void **)&L.b, sizeof(double) * ((size_t)L.own + 1), st));
    MGCG_CUDA(pool_alloc((void **)&L.dinv, sizeof(double) * ((size_t)L.own + 1), st));
    return MGCG_OK;
}


// ---------------------------------------------------------------------------------------------------------
// Build level l+1's owned rows (and level l's P and R) from distributed level l.  Every rank fetches two halo widths of
// matrix rows from each neighbour once, then runs the single-GPU hierarchy kernels on its extended block: prolongator
// rows for [own +- 2*halo], A*P for [own +- halo], R*(A*P) for the aggregates it owns.  No other setup traffic.
static int dist_build_next(mgcg_ctx *c, int l, DevCsr *Ac_out)
{
    DistState *D = c->dist;
    NcclApi *a = D->api;
    cudaStream_t st = c->stream;
    DistLevel &L = D->L[l];
    const LevelGeom &g = L.geom;
    const int W = D->world, rk = D->rank, B = L.halo, own = L.own;
    const int e_lo = rk > 0 ? 2 * B : 0, e_hi = rk < W - 1 ? 2 * B : 0;
    const int lo1 = L.lo, hi1 = L.hi;
    PhaseTrace tr(st);
    if (rk != 0) tr.on = false;
    // ---- fetch the ghost rows: row pointers first (fixed size), then the entries
    int *rp_lo = nullptr, *rp_hi = nullptr;
    MGCG_CUDA(pool_alloc((void **)&rp_lo, sizeof(int) * ((size_t)e_lo + 1), c->stream));
    MGCG_CUDA(pool_alloc((void **)&rp_hi, sizeof(int) * ((size_t)e_hi + 1), c->stream));
    MGCG_CUDA(cudaMemsetAsync(rp_lo, 0, sizeof(int) * ((size_t)e_lo + 1), st));
    MGCG_CUDA(cudaMemsetAsync(rp_hi, 0, sizeof(int) * ((size_t)e_hi + 1), st));
    MGCG_NCCL(a, GroupStart());
    if (rk > 0) {
        MGCG_NCCL(a, Send(L.A.rowptr, (size_t)2 * B + 1, ncclInt, rk - 1, D->comm, st));
        MGCG_NCCL(a, Recv(rp_lo, (size_t)2 * B + 1, ncclInt, rk - 1, D->comm, st));
    }
    if (rk < W - 1) {
        MGCG_NCCL(a, Send(L.A.rowptr + own - 2 * B, (size_t)2 * B + 1, ncclInt, rk + 1, D->comm, st));
        MGCG_NCCL(a, Recv(rp_hi, (size_t)2 * B + 1, ncclInt, rk + 1, D->comm, st));
    }
    MGCG_NCCL(a, GroupEnd());
    int h[8] = {0};
    MGCG_CUDA(cudaMemcpyAsync(&h[0], L.A.rowptr, sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaMemcpyAsync(&h[1], L.A.rowptr + (rk > 0 ? 2 * B : 0), sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaMemcpyAsync(&h[2], L.A.rowptr + (rk < W - 1 ? own - 2 * B : own), sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaMemcpyAsync(&h[3], L.A.rowptr + own, sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaMemcpyAsync(&h[4], rp_lo, sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaMemcpyAsync(&h[5], rp_lo + e_lo, sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaMemcpyAsync(&h[6], rp_hi, sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaMemcpyAsync(&h[7], rp_hi + e_hi, sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    const int nnz_own = h[3] - h[0], nnz_lo = h[5] - h[4], nnz_hi = h[7] - h[6];
    DevCsr E;   // the extended block
    E.rows = e_lo + own + e_hi; E.cols = L.n_glob; E.nnz = nnz_lo + nnz_own + nnz_hi;
    MGCG_CUDA(pool_alloc((void **)&E.rowptr, sizeof(int) * ((size_t)E.rows + 1), c->stream));
    MGCG_CUDA(pool_alloc((void **)&E.col, sizeof(int) * ((size_t)E.nnz + 1), c->stream));
    MGCG_CUDA(pool_alloc((void **)&E.val, sizeof(double) * ((size_t)E.nnz + 1), c->stream));
    MGCG_NCCL(a, GroupStart());
    if (rk > 0) {
        MGCG_NCCL(a, Send(L.A.col + h[0], (size_t)(h[1] - h[0]), ncclInt, rk - 1, D->comm, st));
        MGCG_NCCL(a, Send(L.A.val + h[0], (size_t)(h[1] - h[0]), ncclDouble, rk - 1, D->comm, st));
        MGCG_NCCL(a, Recv(E.col, (size_t)nnz_lo, ncclInt, rk - 1, D->comm, st));
        MGCG_NCCL(a, Recv(E.val, (size_t)nnz_lo, ncclDouble, rk - 1, D->comm, st));
    }
    if (rk < W - 1) {
        MGCG_NCCL(a, Send(L.A.col + h[2], (size_t)(h[3] - h[2]), ncclInt, rk + 1, D->comm, st));
        MGCG_NCCL(a, Send(L.A.val + h[2], (size_t)(h[3] - h[2]), ncclDouble, rk + 1, D->comm, st));
        MGCG_NCCL(a, Recv(E.col + nnz_lo + nnz_own, (size_t)nnz_hi, ncclInt, rk + 1, D->comm, st));
        MGCG_NCCL(a, Recv(E.val + nnz_lo + nnz_own, (size_t)nnz_hi, ncclDouble, rk + 1, D->comm, st));
    }
    MGCG_NCCL(a, GroupEnd());
    MGCG_CUDA(cudaMemcpyAsync(E.col + nnz_lo, L.A.col + h[0], sizeof(int) * (size_t)nnz_own, cudaMemcpyDeviceToDevice, st));
    MGCG_CUDA(cudaMemcpyAsync(E.val + nnz_lo, L.A.val + h[0], sizeof(double) * (size_t)nnz_own, cudaMemcpyDeviceToDevice, st));
    k_ext_rowptr<<<gridn(E.rows + 1), 256, 0, st>>>(e_lo, own, e_hi, rp_lo, L.A.rowptr, rp_hi, nnz_lo, nnz_own, E.rowptr);
    MGCG_LAUNCH_CHECK();
    tr.mark(l, "d:ghost rows");
    const int e0 = L.off - e_lo;                       // global index of extended row 0
    // ---- aggregate ids for every column the extended rows can reference
    int g0 = e0 - B; if (g0 < 0) g0 = 0;
    int g1 = e0 + E.rows + B; if (g1 > L.n_glob) g1 = L.n_glob;
    int *agg = nullptr;
    MGCG_CUDA(pool_alloc((void **)&agg, sizeof(int) * (size_t)(g1 - g0), c->stream));
    k_fill_agg<<<gridn(g1 - g0), 256, 0, st>>>(g, g0, g1 - g0, agg);
    MGCG_LAUNCH_CHECK();
    const int *agg_sh = agg - g0;                      // indexed by global row id
    double *dinv_ext = nullptr;
    MGCG_CUDA(pool_alloc((void **)&dinv_ext, sizeof(double) * ((size_t)E.rows + 1), c->stream));
    MGCG_CHECK(extract_dinv(c, E.view(), e0, dinv_ext));
    // ---- prolongator rows of the whole extended block (global coarse columns)
    const mgcg_options &o = c->opt;
    const bool smooth_here = o.mg_smoothed && (o.mg_smoothed_levels <= 0 || l < o.mg_smoothed_levels);
    const int nc_glob = g.coff[W], coff = g.coff[rk], cown = g.coff[rk + 1] - g.coff[rk];
    MGCG_CHECK(build_prolongator(c, E.view(), e0, agg_sh, dinv_ext, smooth_here, o.mg_prolong_omega, nc_glob, &L.Pext));
    tr.mark(l, "d:P ext");
    // ---- A*P for rows [own +- halo]
    Csr Av = E.view();
    Av.rows = lo1 + own + hi1; Av.rowptr = E.rowptr + (e_lo - lo1);
    DevCsr AP;
    MGCG_CHECK(spgemm(c, Av, L.Pext.view(), &AP, e0));
    tr.mark(l, "d:A*P");
    // ---- R = (P rows [own +- halo])^T restricted to the owned aggregates
    Csr Pv = L.Pext.view();
    Pv.rows = Av.rows; Pv.rowptr = L.Pext.rowptr + (e_lo - lo1);
    int pr[2] = {0, 0};
    MGCG_CUDA(cudaMemcpyAsync(&pr[0], Pv.rowptr, sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaMemcpyAsync(&pr[1], Pv.rowptr + Pv.rows, sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    const size_t cnt = (size_t)(pr[1] - pr[0]);
    unsigned long long *keys = nullptr; double *vals = nullptr;
    MGCG_CUDA(pool_alloc((void **)&keys, sizeof(unsigned long long) * (cnt + 1), c->stream));
    MGCG_CUDA(pool_alloc((void **)&vals, sizeof(double) * (cnt + 1), c->stream));
    k_transpose_expand_filtered<<<gridn(Pv.rows), 256, 0, st>>>(Pv, coff, cown, keys, vals);
    MGCG_LAUNCH_CHECK();
    MGCG_CHECK(coo_to_csr_sum(c, cown + 1, Pv.rows, cnt, keys, vals, &L.R));
    pool_release(keys); pool_release(vals);
    int rn = 0;
    MGCG_CHECK(fetch_int(c, &rn, L.R.rowptr + cown));
    L.R.rows = cown; L.R.nnz = rn;                      // drop the dump row
    tr.mark(l, "d:R");
    // ---- Galerkin rows of the owned aggregates
    MGCG_CHECK(spgemm(c, L.R.view(), AP.view(), Ac_out, 0));
    Ac_out->cols = nc_glob;
    AP.free_all();
    tr.mark(l, "d:R*(AP)");
    // ---- solve-time views
    L.P = L.Pext.view();
    L.P.rows = own; L.P.rowptr = L.Pext.rowptr + e_lo;
    int pn[2] = {0, 0};
    MGCG_CHECK(fetch_int(c, &pn[0], L.P.rowptr));
    MGCG_CHECK(fetch_int(c, &pn[1], L.P.rowptr + own));
    L.P.nnz = pn[1] - pn[0];
    L.planP = spmv_make_plan(c, own, L.P.nnz, MGCG_SPMV_AUTO);
    L.P.nnz = L.Pext.nnz;   // the view addresses val/col with the parent's absolute offsets
    L.planR = spmv_make_plan(c, L.R.rows, L.R.nnz, MGCG_SPMV_AUTO);
    MGCG_CUDA(cudaStreamSynchronize(st));
    pool_release(rp_lo); pool_release(rp_hi); pool_release(agg); pool_release(dinv_ext);
    E.free_all();
    tr.mark(l, "d:finish");
    return MGCG_OK;
}


// all-gather with per-rank counts: grouped point-to-point (every rank's block lands at offs[r] of `full`; the caller has
// already placed its own block there)
static int allgather_v(mgcg_ctx *c, void *full, const int *offs, ncclDataType_t type, size_t elem)
{
    DistState *D = c->dist;
    NcclApi *a = D->api;
    if (D->world == 1) return MGCG_OK;
    char *base = (char *)full;
    MGCG_NCCL(a, GroupStart());
    for (int q = 0; q < D->world; ++q) {
        if (q == D->rank) continue;
        MGCG_NCCL(a, Send(base + elem * (size_t)offs[D->rank], (size_t)(offs[D->rank + 1] - offs[D->rank]), type, q, D->comm, c->stream));
        MGCG_NCCL(a, Recv(base + elem * (size_t)offs[q], (size_t)(offs[q + 1] - offs[q]), type, q, D->comm, c->stream));
    }
    MGCG_NCCL(a, GroupEnd());
    return MGCG_OK;
}

// Replicate level `lev` (owned rows A, global columns) on every rank and build the rest of the hierarchy there with the
// single-GPU code in a private context that shares this context's stream and CG scalars.
static int dist_replicate(mgcg_ctx *c, int lev, const Csr &A, const int *offs, int nx, int ny, int nz_total)
{
    DistState *D = c->dist;
    cudaStream_t st = c->stream;
    const int W = D->world, rk = D->rank;
    const int n = offs[W], own = offs[rk + 1] - offs[rk];
    PhaseTrace tr(st);
    if (rk != 0) tr.on = false;
    // non-zero counts of every rank
    int my[2] = {0, 0};
    MGCG_CHECK(fetch_int(c, &my[0], A.rowptr));
    MGCG_CHECK(fetch_int(c, &my[1], A.rowptr + own));
    int *d_cnt = nullptr;
    MGCG_CUDA(pool_alloc((void **)&d_cnt, sizeof(int) * (size_t)(W + 1), c->stream));
    const int mynnz = my[1] - my[0];
    MGCG_CUDA(cudaMemcpyAsync(d_cnt + W, &mynnz, sizeof(int), cudaMemcpyHostToDevice, st));
    MGCG_NCCL(D->api, AllGather(d_cnt + W, d_cnt, 1, ncclInt, D->comm, st));
    int cnts[kMaxRanks], zoff[kMaxRanks + 1];
    MGCG_CUDA(cudaMemcpyAsync(cnts, d_cnt, sizeof(int) * (size_t)W, cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    pool_release(d_cnt);
    zoff[0] = 0;
    long long tot = 0;
    for (int r = 0; r < W; ++r) { tot += cnts[r]; zoff[r + 1] = (int)tot; }
    if (tot >= (1ll << 31)) { set_error("dist_replicate: replicated level exceeds int32 non-zeros"); return MGCG_ERR_UNSUPPORTED; }
    DevCsr &F = D->coarse_A;
    F.rows = n; F.cols = n; F.nnz = (int)tot;
    MGCG_CUDA(pool_alloc((void **)&F.rowptr, sizeof(int) * ((size_t)n + 1), c->stream));
    MGCG_CUDA(pool_alloc((void **)&F.col, sizeof(int) * ((size_t)tot + 1), c->stream));
    MGCG_CUDA(pool_alloc((void **)&F.val, sizeof(double) * ((size_t)tot + 1), c->stream));
    // own block in place: row pointers rebased to the global non-zero numbering
    MGCG_CUDA(cudaMemcpyAsync(F.rowptr + offs[rk], A.rowptr, sizeof(int) * (size_t)own, cudaMemcpyDeviceToDevice, st));
    if (own > 0) { k_add_int<<<gridn(own), 256, 0, st>>>(own, F.rowptr + offs[rk], zoff[rk] - my[0]); MGCG_LAUNCH_CHECK(); }
    MGCG_CUDA(cudaMemcpyAsync(F.col + zoff[rk], A.col + my[0], sizeof(int) * (size_t)mynnz, cudaMemcpyDeviceToDevice, st));
    MGCG_CUDA(cudaMemcpyAsync(F.val + zoff[rk], A.val + my[0], sizeof(double) * (size_t)mynnz, cudaMemcpyDeviceToDevice, st));
    MGCG_CUDA(cudaMemcpyAsync(F.rowptr + n, &F.nnz, sizeof(int), cudaMemcpyHostToDevice, st));
    MGCG_CHECK(allgather_v(c, F.rowptr, offs, ncclInt, sizeof(int)));
    MGCG_CHECK(allgather_v(c, F.col, zoff, ncclInt, sizeof(int)));
    MGCG_CHECK(allgather_v(c, F.val, zoff, ncclDouble, sizeof(double)));
    MGCG_CUDA(cudaStreamSynchronize(st));
    tr.mark(lev, "d:replicate A");
    // right-hand side buffer of the replicated level
    D->coarse_n = n;
    for (int r = 0; r <= W; ++r) D->coarse_offs[r] = offs[r];
    MGCG_CUDA(pool_alloc((void **)&D->coarse_b, sizeof(double) * ((size_t)n + 1), c->stream));
    // the private context
    if (D->coarse_spare) { D->coarse = D->coarse_spare; D->coarse_spare = nullptr; }
    else D->coarse = mgcg_create(c->device);
    if (!D->coarse) return MGCG_ERR_CUDA;
    MGCG_CHECK(mgcg_set_stream(D->coarse, (void *)st));
    D->coarse->child = true;        // its cached blocks live under the parent's stream: the parent trims them
    mgcg_options o = c->opt;
    o.precond = MGCG_PRECOND_MG;
    o.nx = nx; o.ny = ny; o.nz = nz_total;
    // the sharded levels above aggregate without a grid by blocks of 8 rows of each rank's block (agg_of): keep the same rule
    // in the replicated tail so that the sharded hierarchy equals the single-GPU one built with MGCG_AGG_BLOCKS
    if (o.mg_aggregation == MGCG_AGG_AUTO) o.mg_aggregation = MGCG_AGG_BLOCKS;
    if (lev > 0) {
        if (o.mg_aggregate_size_coarse > 0) o.mg_aggregate_size = o.mg_aggregate_size_coarse;
        if (o.mg_smoothed && o.mg_smoothed_levels > 0) {
            o.mg_smoothed_levels -= lev;
            if (o.mg_smoothed_levels <= 0) { o.mg_smoothed = 0; o.mg_smoothed_levels = 0; }
        }
        if (o.mg_max_levels > 0) { o.mg_max_levels -= lev; if (o.mg_max_levels < 1) o.mg_max_levels = 1; }
    }
    MGCG_CHECK(mgcg_setup(D->coarse, n, F.nnz, F.val, F.rowptr, F.col, &o));
    tr.mark(lev, "d:tail setup");
    return MGCG_OK;
}


// ---------------------------------------------------------------------------------------------------------
static void dist_graph_free(DistState *D)
{
    if (D->graph_exec) { cudaGraphExecDestroy(D->graph_exec); D->graph_exec = nullptr; }
    if (D->graph) { cudaGraphDestroy(D->graph); D->graph = nullptr; }
    D->graph_b = nullptr; D->graph_x = nullptr;
}

// Choose the loop's transport once the hierarchy (hence every halo width and the gathered level's size) is known.
// dist_transport: 0 = peer memory when every rank can map every other (one NVLink/NVSwitch node), else NCCL;
// 1 = NCCL; 2 = peer memory or fail.  MGCG_DIST_TRANSPORT=nccl|peer overrides 0.
static int dist_open_transport(mgcg_ctx *c)
{
    DistState *D = c->dist;
    int want = c->opt.dist_transport;
    if (want == MGCG_TRANSPORT_AUTO) {
        const char *e = getenv("MGCG_DIST_TRANSPORT");
        if (e && !strcmp(e, "nccl")) want = MGCG_TRANSPORT_NCCL;
        else if (e && !strcmp(e, "peer")) want = MGCG_TRANSPORT_PEER;
    }
    if (D->world < 2 || want == MGCG_TRANSPORT_NCCL) { peer_close(c, D->world >= 2); return MGCG_OK; }
    int halo_cap = 0;
    for (size_t l = 0; l < D->L.size(); ++l) if (D->L[l].halo > halo_cap && D->L[l].halo < 0x3fffffff) halo_cap = D->L[l].halo;
    // fused halo exchange (default; MGCG_DIST_FUSED=0 keeps the separate exchange kernels): the [lo|own|hi] vectors move into the arena
    const char *ef = getenv("MGCG_DIST_FUSED");
    const bool want_fused = !(ef && ef[0] == '0');
    size_t vec_doubles = 0;
    auto ext_doubles = [](const DistLevel &L) { return (((size_t)L.lo + L.own + L.hi + 2) * sizeof(double) + 255) / 256 * 32; };
    if (want_fused) {
        vec_doubles += 2 * ext_doubles(D->L[0]);
        for (size_t l = 0; l < D->L.size(); ++l) vec_doubles += 3 * ext_doubles(D->L[l]);
    }
    MGCG_CHECK(peer_open(c, halo_cap, D->coarse ? D->coarse_n : 0, vec_doubles));
    if (!D->peer.active && want == MGCG_TRANSPORT_PEER) {
        set_error("mgcg_setup_dist: dist_transport = peer, but the ranks cannot map each other's memory (world %d, limit %d)", D->world, kPeerMax);
        return MGCG_ERR_UNSUPPORTED;
    }
    if (D->peer.active && D->peer.fused) {
        // same order on every rank: p, r, then xa / xb / r of every distributed level
        auto rehome = [&](double **v, const DistLevel &L) -> int {
            double *a = peer_vec_alloc(c, L.lo, L.own, L.hi, L.halo);
            if (!a) { set_error("mgcg_setup_dist: the peer arena's vector region is full"); return MGCG_ERR_INVALID; }
            if (*v) pool_release(*v);
            *v = a;
            return MGCG_OK;
        };
        DistLevel &L0 = D->L[0];
        MGCG_CHECK(rehome(&D->p, L0));
        MGCG_CHECK(rehome(&D->r, L0));
        for (size_t l = 0; l < D->L.size(); ++l) {
            DistLevel &L = D->L[l];
            MGCG_CHECK(rehome(&L.xa, L)); MGCG_CHECK(rehome(&L.xb, L)); MGCG_CHECK(rehome(&L.r, L));
        }
        D->vec_in_arena = true;
        MGCG_CHECK(peer_vec_publish(c));
    }
    return MGCG_OK;
}

static int dist_setup(mgcg_ctx *c, int global_count, int row_offset, int local_count, int local_nnz,
                      const double *val, const int *rowptr, const int *col)
{
    DistState *D = c->dist;
    cudaStream_t st = c->stream;
    const int W = D->world, rk = D->rank;
    const mgcg_options &o = c->opt;
    PhaseTrace tr(st);
    if (rk != 0) tr.on = false;
    dist_graph_free(D);
    D->peer.active = false;          // setup traffic goes over NCCL; the arena is kept for dist_open_transport to reuse
    dist_free_levels(D);
    set_alloc_stream(st);
    pool_reserve((size_t)(2.6 * 12.0 * (double)local_nnz + 240.0 * (double)local_count) + ((size_t)96 << 20));
    tr.mark(-1, "d:free old");
    // ---- the partition: all-gather of the local row counts (ConjugateGradientParallelGpu.cs:271-277 is the caller's rule)
    int offs[kMaxRanks + 1];
    {
        int *d = nullptr;
        MGCG_CUDA(pool_alloc((void **)&d, sizeof(int) * (size_t)(W + 1), c->stream));
        MGCG_CUDA(cudaMemcpyAsync(d + W, &local_count, sizeof(int), cudaMemcpyHostToDevice, st));
        MGCG_NCCL(D->api, AllGather(d + W, d, 1, ncclInt, D->comm, st));
        int cnt[kMaxRanks];
        MGCG_CUDA(cudaMemcpyAsync(cnt, d, sizeof(int) * (size_t)W, cudaMemcpyDeviceToHost, st));
        MGCG_CUDA(cudaStreamSynchronize(st));
        pool_release(d);
        offs[0] = 0;
        for (int r = 0; r < W; ++r) offs[r + 1] = offs[r] + cnt[r];
    }
    if (offs[W] != global_count || offs[rk] != row_offset) {
        set_error("mgcg_setup_dist: the ranks' blocks do not tile [0,%d): rank %d has offset %d, expected %d", global_count, rk, row_offset, offs[rk]);
        return MGCG_ERR_INVALID;
    }
    D->n_glob = global_count; D->off = row_offset; D->own = local_count;
    c->A.rows = local_count; c->A.cols = global_count; c->A.nnz = local_nnz;
    c->A.val = val; c->A.rowptr = rowptr; c->A.col = col;
    c->planA = spmv_make_plan(c, local_count, local_nnz, o.spmv_kernel);
    MGCG_CHECK(spmv_code_build(c, &c->code0, &c->A, &c->planA));     // coded columns: the distance col - (local row) is as regular on a shard

    int nx = 0, ny = 0;
    if (o.nx > 0 && o.ny > 0 && o.nz > 0 && (long long)o.nx * o.ny * o.nz == global_count) { nx = o.nx; ny = o.ny; }
    const int max_levels = o.mg_max_levels > 0 ? o.mg_max_levels : 12;
    const int replicate_below = o.dist_replicate_below > 0 ? o.dist_replicate_below : (1 << 18);

    D->L.emplace_back();
    {
        DistLevel &L = D->L[0];
        L.n_glob = global_count; L.own = local_count; L.off = row_offset; L.A = c->A; L.planA = c->planA;
        MGCG_CHECK(level_halfband(c, L.A, L.off, &L.halo));
        for (int r = 0; r < W; ++r)
            if (W > 1 && offs[r + 1] - offs[r] < L.halo) {
                set_error("mgcg_setup_dist: half bandwidth %d exceeds rank %d's block of %d rows (only neighbouring ranks are exchanged with, "
                          "like the reference's halo rule, ConjugateGradientParallelGpu.cs:397-398)", L.halo, r, offs[r + 1] - offs[r]);
                return MGCG_ERR_UNSUPPORTED;
            }
        L.lo = rk > 0 ? L.halo : 0; L.hi = rk < W - 1 ? L.halo : 0;
        MGCG_CHECK(alloc_level_vectors_dist(L, true, st));
        MGCG_CHECK(extract_dinv(c, L.A, L.off, L.dinv));
        if (o.precond == MGCG_PRECOND_MG) MGCG_CHECK(dist_level_chebyshev(c, L));
        const size_t e = (size_t)L.lo + L.own + L.hi + 1;
        MGCG_CUDA(pool_alloc((void **)&D->p, sizeof(double) * e, c->stream));
        MGCG_CUDA(pool_alloc((void **)&D->r, sizeof(double) * e, c->stream));
        MGCG_CUDA(pool_alloc((void **)&D->Ap, sizeof(double) * ((size_t)L.own + 1), c->stream));
        MGCG_CUDA(cudaMemsetAsync(D->p, 0, sizeof(double) * e, st));
        MGCG_CUDA(cudaMemsetAsync(D->r, 0, sizeof(double) * e, st));
    }
    D->first_replicated = 0;
    tr.mark(0, "d:level0 init");
    if (o.precond != MGCG_PRECOND_MG) return dist_open_transport(c);

    // ---- hierarchy: distributed levels while they are large and their halo fits, then the replicated tail
    int cur_offs[kMaxRanks + 1];
    memcpy(cur_offs, offs, sizeof(int) * (size_t)(W + 1));
    int cur_nx = nx, cur_ny = ny;
    for (int l = 0;; ++l) {
        DistLevel &L = D->L[l];
        int min_own = 0x7fffffff;
        for (int r = 0; r < W; ++r) min_own = cur_offs[r + 1] - cur_offs[r] < min_own ? cur_offs[r + 1] - cur_offs[r] : min_own;
        bool distribute = W > 1 && L.n_glob >= replicate_below && (long long)min_own >= 2ll * L.halo && l + 1 < max_levels;
        int a = (l == 0 || o.mg_aggregate_size_coarse <= 0) ? o.mg_aggregate_size : o.mg_aggregate_size_coarse;
        if (a < 2) a = 2;
        if (distribute) {
            make_geom(L, W, cur_offs, cur_nx, cur_ny, a);
            if (L.geom.coff[W] >= L.n_glob) distribute = false;
            for (int r = 0; r < W; ++r) if (L.geom.coff[r + 1] == L.geom.coff[r]) distribute = false;
        }
        if (!distribute) {
            bool structured = cur_nx > 0 && cur_ny > 0;
            if (structured) for (int r = 0; r < W; ++r) if ((cur_offs[r + 1] - cur_offs[r]) % (cur_nx * cur_ny) != 0) structured = false;
            const int nzt = structured ? L.n_glob / (cur_nx * cur_ny) : 0;
            MGCG_CHECK(dist_replicate(c, l, L.A, cur_offs, structured ? cur_nx : 0, structured ? cur_ny : 0, nzt));
            D->first_replicated = l;
            if (l > 0) {   // level l lives in the private context only: drop its distributed shell (the matrix was copied)
                D->L[l].Aown.free_all();
                if (D->L[l].dinv) pool_release(D->L[l].dinv);
                if (D->L[l].xa) pool_release(D->L[l].xa);
                if (D->L[l].xb) pool_release(D->L[l].xb);
                if (D->L[l].r) pool_release(D->L[l].r);
                if (D->L[l].d) pool_release(D->L[l].d);
                if (D->L[l].b) pool_release(D->L[l].b);
                D->L.pop_back();
            }
            break;
        }
        DevCsr Ac;
        MGCG_CHECK(dist_build_next(c, l, &Ac));
        const LevelGeom g = D->L[l].geom;
        D->L.emplace_back();
        DistLevel &N = D->L[l + 1];
        N.Aown = Ac; N.A = Ac.view();
        N.n_glob = g.coff[W]; N.own = g.coff[rk + 1] - g.coff[rk]; N.off = g.coff[rk];
        N.planA = spmv_make_plan(c, N.own, N.A.nnz, MGCG_SPMV_AUTO);
        MGCG_CHECK(level_halfband(c, N.A, N.off, &N.halo));
        int nmin = 0x7fffffff;
        for (int r = 0; r < W; ++r) nmin = g.coff[r + 1] - g.coff[r] < nmin ? g.coff[r + 1] - g.coff[r] : nmin;
        const bool halo_fits = nmin >= N.halo;
        N.lo = (rk > 0 && halo_fits) ? N.halo : 0; N.hi = (rk < W - 1 && halo_fits) ? N.halo : 0;
        if (!halo_fits) N.halo = 0x3fffffff;   // forces replication of this level on the next pass
        MGCG_CHECK(alloc_level_vectors_dist(N, false, st));
        MGCG_CHECK(extract_dinv(c, N.A, N.off, N.dinv));
        MGCG_CHECK(dist_level_chebyshev(c, N));
        memcpy(cur_offs, g.coff, sizeof(int) * (size_t)(W + 1));
        if (g.structured) { cur_nx = g.cx; cur_ny = g.cy; } else { cur_nx = cur_ny = 0; }
    }
    // the private context runs on this context's scalars (its kernels are gated by the same `done` flag)
    D->coarse_scalars_orig = D->coarse->d_scalars;
    D->coarse->d_scalars = c->d_scalars;
    tr.mark(-1, "d:levels");
    int s_ = dist_open_transport(c);
    tr.mark(-1, "d:transport");
    return s_;
}


// ---------------------------------------------------------------------------------------------------------
// solve: per-rank partial sums -> all-gather -> every rank adds them in rank order (deterministic, identical on all ranks;
// the reference sums its per-device partial dots on the host in device order, ConjugateGradientParallelGpu.cs:463,499,525)
__global__ void k_dist_finalize(int mode, int k, int world, const double *__restrict__ all, Scalars *sc, bool gated)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (gated && sc->done) return;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
    for (int r = 0; r < world; ++r) { s0 += all[r * k]; if (k > 1) s1 += all[r * k + 1]; if (k > 2) s2 += all[r * k + 2]; }
    dist_finalize(mode, s0, s1, s2, sc);
}

static int dist_reduce(mgcg_ctx *c, int mode, int k, bool gated)
{
    DistState *D = c->dist;
    D->allgathers++;
    if (D->peer.active) return peer_reduce(c, mode, k, gated);
    MGCG_NCCL(D->api, AllGather(D->partial_send, D->partial_all, (size_t)k, ncclDouble, D->comm, c->stream));
    k_dist_finalize<<<1, 32, 0, c->stream>>>(mode, k, D->world, D->partial_all, c->d_scalars, gated);
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

static inline const double *shifted(const double *ext, int off, int lo) { return ext - ((ptrdiff_t)off - lo); }
// x of a product = the extended vector [lo ghosts | own rows | hi ghosts] seen through global column indices; its readable range
static inline void set_shifted_x(SpmvArgs &a, const double *ext, int off, int lo, int own, int hi)
{
    a.x = shifted(ext, off, lo); a.x_lo = off - lo; a.x_hi = off + own + hi;
}

// ---- fused halo exchange (halo.cuh): with the level vectors in the peer arena, the kernel that WRITES a [lo|own|hi] vector stores
// its boundary rows into the neighbours' ghost rows and publishes the exchange, and the kernel that READS ghost rows waits for
// the neighbours' flags — dist_halo_exchange's separate launch disappears from the loop.  Why the direct ghost stores cannot
// overwrite rows a neighbour is still reading: between two stores into the same buffer there is always either a scalar reduction
// or the coarse all-gather (every rank waits for every rank, after its readers in stream order), or the storing kernel has itself
// waited for a flag the neighbour published at the END of a kernel that follows its last reader of that buffer (the sweeps of a
// level alternate xa / xb, and every sweep both waits and publishes).
static inline bool dist_fused(const mgcg_ctx *c) { return c->dist->peer.active && c->dist->peer.fused; }
// producer descriptor of ext vector v (counts as one exchange); mir.ctl stays null when the exchange is not fused
static inline Mirror mirror_for(mgcg_ctx *c, const double *v_ext)
{
    Mirror m;
    if (dist_fused(c) && peer_mirror_of(c, v_ext, &m)) c->dist->halo_exchanges++;
    return m;
}
static inline void wait_for_ghosts(mgcg_ctx *c, SpmvArgs &a) { if (dist_fused(c)) peer_wait_of(c, &a.gw); }
// explicit exchange of v unless its producer mirrored it
static inline int halo_unless_fused(mgcg_ctx *c, double *v, int own, int lo, int hi, int halo)
{
    Mirror m;
    if (dist_fused(c) && peer_mirror_of(c, v, &m)) return MGCG_OK;
    return dist_halo_exchange(c, v, own, lo, hi, halo);
}

static int dist_gather_coarse_b(mgcg_ctx *c)
{
    DistState *D = c->dist;
    if (D->peer.active) return peer_allgather(c, D->coarse_b, D->coarse_offs);
    return allgather_v(c, D->coarse_b, D->coarse_offs, ncclDouble, sizeof(double));
}

// Chebyshev smoother on a distributed level: the eigenvalue bound is the maximum over every rank's rows
static int dist_level_chebyshev(mgcg_ctx *c, DistLevel &L)
{
    DistState *D = c->dist;
    if (c->opt.mg_smoother != MGCG_SMOOTHER_CHEBYSHEV) return MGCG_OK;
    L.cheb = c->opt.mg_cheby_degree;
    double mine = 0.0;
    MGCG_CHECK(gershgorin_lmax(c, L.A, L.dinv, &mine));
    double *d = nullptr;
    MGCG_CUDA(pool_alloc((void **)&d, 2 * sizeof(double), c->stream));
    MGCG_CUDA(cudaMemcpyAsync(d, &mine, sizeof(double), cudaMemcpyHostToDevice, c->stream));
    MGCG_NCCL(D->api, AllReduce(d, d + 1, 1, ncclDouble, ncclMax, D->comm, c->stream));
    MGCG_CUDA(cudaMemcpyAsync(&L.lmax, d + 1, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    MGCG_CUDA(cudaStreamSynchronize(c->stream));
    pool_release(d);
    MGCG_CUDA(pool_alloc((void **)&L.d, sizeof(double) * ((size_t)L.own + 1), c->stream));
    return MGCG_OK;
}
static double dist_first_omega(mgcg_ctx *c, const DistLevel &L)
{
    if (!L.cheb) return c->opt.mg_omega;
    double c1, c2;
    cheby_coeffs(L.lmax, c->opt.mg_cheby_ratio, 0, &c1, &c2);
    return c2;
}

// one Chebyshev smoothing step (L.cheb products, a halo exchange before each); see mg_cycle.cu cheby_smooth
static int dist_cheby_smooth(mgcg_ctx *c, DistLevel &L, bool skip_first, double *dot_out, bool ghosts_needed = true)
{
    for (int k = skip_first ? 1 : 0; k < L.cheb; ++k) {
        double c1, c2;
        cheby_coeffs(L.lmax, c->opt.mg_cheby_ratio, k, &c1, &c2);
        double *other = (L.xcur == L.xa) ? L.xb : L.xa;
        const bool last = dot_out && k == L.cheb - 1;
        MGCG_CHECK(halo_unless_fused(c, L.xcur, L.own, L.lo, L.hi, L.halo));
        SpmvArgs a;
        a.A = L.A; set_shifted_x(a, L.xcur, L.off, L.lo, L.own, L.hi); a.xrow = L.xcur + L.lo; a.y = other + L.lo; a.b = L.b; a.dinv = L.dinv;
        wait_for_ghosts(c, a);
        if (ghosts_needed || k + 1 < L.cheb) a.mir = mirror_for(c, other);
        a.alpha = c1; a.omega = c2; a.d_in = (skip_first && k == 1) ? L.xcur + L.lo : L.d; a.d_out = L.d;
        a.gated = true; a.fin = FIN_STORE; a.out = dot_out;
        MGCG_CHECK(spmv_launch(c, L.planA, last ? EPI_CHEBY_DOT : EPI_CHEBY, a));
        L.xcur = other;
    }
    return MGCG_OK;
}

static int dist_jacobi_sweep(mgcg_ctx *c, DistLevel &L, double omega, double *dot_out, bool ghosts_needed = true)
{
    double *other = (L.xcur == L.xa) ? L.xb : L.xa;
    MGCG_CHECK(halo_unless_fused(c, L.xcur, L.own, L.lo, L.hi, L.halo));
    SpmvArgs a;
    a.A = L.A; set_shifted_x(a, L.xcur, L.off, L.lo, L.own, L.hi); a.xrow = L.xcur + L.lo; a.y = other + L.lo; a.b = L.b; a.dinv = L.dinv;
    wait_for_ghosts(c, a);
    if (ghosts_needed) a.mir = mirror_for(c, other);
    a.omega = omega; a.gated = true; a.fin = FIN_STORE; a.out = dot_out;
    MGCG_CHECK(spmv_launch(c, L.planA, dot_out ? EPI_JACOBI_DOT : EPI_JACOBI, a));
    L.xcur = other;
    return MGCG_OK;
}

// z = M r for the sharded hierarchy.  r_ext is the [lo|own|hi] CG residual; *z_own receives the own-rows pointer of z.
// rz_out != nullptr: the local part of r.z is left there (fused into the last post-smoothing sweep when there is one).
static int dist_vcycle(mgcg_ctx *c, double *r_ext, double **z_own, bool fuse_first_done, double *rz_out)
{
    DistState *D = c->dist;
    const int Ld = D->first_replicated;
    const double omega = c->opt.mg_omega;
    const int pre = c->opt.mg_pre_sweeps, post = c->opt.mg_post_sweeps;
    DistLevel &L0 = D->L[0];
    double *r_own = r_ext + L0.lo;
    if (Ld == 0) {   // the whole preconditioner is replicated: gather r, cycle, take the own slice
        MGCG_CUDA(cudaMemcpyAsync(D->coarse_b + D->coarse_offs[D->rank], r_own, sizeof(double) * (size_t)L0.own, cudaMemcpyDeviceToDevice, c->stream));
        MGCG_CHECK(dist_gather_coarse_b(c));
        double *z = nullptr;
        MGCG_CHECK(mg_vcycle(D->coarse, D->coarse_b, &z, false, -1));
        *z_own = z + L0.off;
        if (rz_out) MGCG_CHECK(vec_dot(c, L0.own, r_own, *z_own, FIN_STORE, rz_out));
        return MGCG_OK;
    }
    L0.b = r_own;
    for (int l = 0; l < Ld; ++l) {
        DistLevel &L = D->L[l];
        L.xcur = L.xa;
        if (!(l == 0 && fuse_first_done)) {
            const Mirror m1 = mirror_for(c, L.xa);
            MGCG_CHECK(mg_jacobi_first(c, L.own, dist_first_omega(c, L), L.dinv, L.b, L.xa + L.lo, &m1));
        }
        if (L.cheb) {
            MGCG_CHECK(dist_cheby_smooth(c, L, true, nullptr));
            for (int s = 1; s < pre; ++s) MGCG_CHECK(dist_cheby_smooth(c, L, false, nullptr));
        } else {
            for (int s = 1; s < pre; ++s) MGCG_CHECK(dist_jacobi_sweep(c, L, omega, nullptr));
        }
        dist_trace(c, l == 0 ? "L0 pre-smooth" : "L1+ pre-smooth");
        MGCG_CHECK(halo_unless_fused(c, L.xcur, L.own, L.lo, L.hi, L.halo));
        dist_trace(c, l == 0 ? "L0 halo(z)" : "L1+ halo(z)");
        SpmvArgs a;
        a.A = L.A; set_shifted_x(a, L.xcur, L.off, L.lo, L.own, L.hi); a.y = L.r + L.lo; a.b = L.b; a.gated = true;
        wait_for_ghosts(c, a); a.mir = mirror_for(c, L.r);
        MGCG_CHECK(spmv_launch(c, L.planA, EPI_RESID, a));
        dist_trace(c, l == 0 ? "L0 residual" : "L1+ residual");
        MGCG_CHECK(halo_unless_fused(c, L.r, L.own, L.lo, L.hi, L.halo));
        dist_trace(c, l == 0 ? "L0 halo(r)" : "L1+ halo(r)");
        SpmvArgs q;
        q.A = L.R.view(); q.x = L.r; q.gated = true;
        wait_for_ghosts(c, q);
        q.y = (l + 1 < Ld) ? D->L[l + 1].b : D->coarse_b + D->coarse_offs[D->rank];
        MGCG_CHECK(spmv_launch(c, L.planR, EPI_PLAIN, q));
        dist_trace(c, l == 0 ? "L0 restrict" : "L1+ restrict");
    }
    MGCG_CHECK(dist_gather_coarse_b(c));
    dist_trace(c, "gather coarse rhs");
    double *zc = nullptr;
    MGCG_CHECK(mg_vcycle(D->coarse, D->coarse_b, &zc, false, -1));
    dist_trace(c, "replicated tail cycle");
    for (int l = Ld - 1; l >= 0; --l) {
        DistLevel &L = D->L[l];
        const double *xn;
        if (l + 1 == Ld) xn = zc;
        else {
            DistLevel &N = D->L[l + 1];
            MGCG_CHECK(halo_unless_fused(c, N.xcur, N.own, N.lo, N.hi, N.halo));
            dist_trace(c, "L1+ halo(xc)");
            xn = shifted(N.xcur, N.off, N.lo);
        }
        SpmvArgs a;
        a.A = L.P; a.x = xn; a.y = L.xcur + L.lo; a.alpha = 1.0; a.beta = 1.0; a.gated = true;
        if (l + 1 < Ld) wait_for_ghosts(c, a);
        if (post > 0 || l > 0) a.mir = mirror_for(c, L.xcur);
        MGCG_CHECK(spmv_launch(c, L.planP, EPI_AXPBY, a));
        dist_trace(c, l == 0 ? "L0 prolong" : "L1+ prolong");
        for (int s = 0; s < post; ++s) {
            const bool last = (l == 0 && s == post - 1 && rz_out);
            // the ghost rows of the cycle's result z are read only by the single-reduce recurrence (w = A z)
            const bool ghosts = !(l == 0 && s == post - 1) || c->opt.cg_variant == MGCG_CG_SINGLE_REDUCE;
            if (L.cheb) MGCG_CHECK(dist_cheby_smooth(c, L, false, last ? rz_out : nullptr, ghosts));
            else MGCG_CHECK(dist_jacobi_sweep(c, L, omega, last ? rz_out : nullptr, ghosts));
            dist_trace(c, l == 0 ? "L0 post-smooth (halo + sweep)" : "L1+ post-smooth (halo + sweep)");
        }
    }
    *z_own = L0.xcur + L0.lo;
    if (post == 0 && rz_out) MGCG_CHECK(vec_dot(c, L0.own, r_own, *z_own, FIN_STORE, rz_out));
    return MGCG_OK;
}


// ---------------------------------------------------------------------------------------------------------
// one CG iteration on the shards (same algebra and order as cg.cu / ConjugateGradientCpu.cs:70-94)
static int dist_enqueue_iteration(mgcg_ctx *c, double *x)
{
    DistState *D = c->dist;
    DistLevel &L0 = D->L[0];
    const int n = L0.own;
    const bool pre = (c->opt.precond == MGCG_PRECOND_MG);
    double *p_own = D->p + L0.lo, *r_own = D->r + L0.lo;
    // halo of p, Ap = A p, local p.Ap                                    (SyncP + Solve1, cs:470-499)
    dist_trace(c, "update p");
    MGCG_CHECK(halo_unless_fused(c, D->p, n, L0.lo, L0.hi, L0.halo));    // fused: pushed by the update of p (by dist_solve before the first iteration)
    dist_trace(c, "halo(p)");
    SpmvArgs a;
    a.A = c->A; set_shifted_x(a, D->p, L0.off, L0.lo, L0.own, L0.hi); a.y = D->Ap; a.w = p_own; a.fin = FIN_STORE; a.out = D->partial_send; a.gated = true;
    wait_for_ghosts(c, a);
    MGCG_CHECK(spmv_launch(c, c->planA, EPI_DOT, a));
    dist_trace(c, "A p + p.Ap");
    MGCG_CHECK(dist_reduce(c, DF_PAP, 1, true));
    dist_trace(c, "reduce p.Ap");
    if (!pre) {
        MGCG_CHECK(cg_update_xr(c, n, p_own, D->Ap, x, r_own, FIN_STORE, false, 0.0, D->partial_send));   // Solve2, cs:505-525
        MGCG_CHECK(dist_reduce(c, DF_RR, 1, true));
        const Mirror mp = mirror_for(c, D->p);
        MGCG_CHECK(cg_update_p(c, n, r_own, p_own, false, 0.0, &mp));                                     // Solve3, cs:548-556
    } else {
        const bool fuse = D->first_replicated > 0;
        const Mirror mz = fuse ? mirror_for(c, L0.xa) : Mirror();
        MGCG_CHECK(cg_update_xr(c, n, p_own, D->Ap, x, r_own, FIN_STORE, false, 0.0, D->partial_send,
                                fuse ? L0.xa + L0.lo : nullptr, fuse ? L0.dinv : nullptr, dist_first_omega(c, L0), &mz));
        dist_trace(c, "update x,r (+z0)");
        double *z = nullptr;
        MGCG_CHECK(dist_vcycle(c, D->r, &z, fuse, D->partial_send + 1));
        MGCG_CHECK(dist_reduce(c, DF_RR_RZ, 2, true));   // stop rule on r.r, beta from r.z: one exchange for both
        dist_trace(c, "reduce r.r, r.z");
        const Mirror mp = mirror_for(c, D->p);
        MGCG_CHECK(cg_update_p(c, n, z, p_own, false, 0.0, &mp));
    }
    return MGCG_OK;
}

// One iteration of the single-reduce recurrence on the shards (cg.cu enqueue_iteration_cgcg): ONE scalar exchange carrying w.u, r.u
// and r.r instead of the classic loop's two.
static int dist_enqueue_iteration_cgcg(mgcg_ctx *c, double *x)
{
    DistState *D = c->dist;
    DistLevel &L0 = D->L[0];
    const int n = L0.own;
    const bool pre = (c->opt.precond == MGCG_PRECOND_MG);
    double *p_own = D->p + L0.lo, *r_own = D->r + L0.lo;
    double *u_ext = D->r;
    if (!pre) {
        MGCG_CHECK(cgcg_update(c, n, r_own, D->Ap, p_own, D->s, x, r_own, nullptr, nullptr, 0.0, D->partial_send + 2));
    } else {
        const bool fuse = D->first_replicated > 0;
        const Mirror mz = fuse ? mirror_for(c, L0.xa) : Mirror();
        MGCG_CHECK(cgcg_update(c, n, D->cgcg_u, D->Ap, p_own, D->s, x, r_own, fuse ? L0.xa + L0.lo : nullptr, fuse ? L0.dinv : nullptr,
                               dist_first_omega(c, L0), D->partial_send + 2, &mz));
        dist_trace(c, "update p,s,x,r (+z0)");
        double *z = nullptr;
        MGCG_CHECK(dist_vcycle(c, D->r, &z, fuse, D->partial_send + 1));
        if (z != D->cgcg_u) { set_error("mgcg_solve_dist: the V-cycle's output buffer moved between iterations"); return MGCG_ERR_INVALID; }
        u_ext = z - L0.lo;
    }
    if (!(pre && D->first_replicated == 0)) {    // (a fully replicated preconditioner returns a slice of a global-length vector)
        // fused: u = the V-cycle's last post-smoothing output was pushed by that sweep; the plain recurrence's u = r was not
        if (pre && D->first_replicated > 0 && c->opt.mg_post_sweeps > 0) MGCG_CHECK(halo_unless_fused(c, u_ext, n, L0.lo, L0.hi, L0.halo));
        else MGCG_CHECK(dist_halo_exchange(c, u_ext, n, L0.lo, L0.hi, L0.halo));
        dist_trace(c, "halo(u)");
    }
    SpmvArgs a;
    a.A = c->A; a.y = D->Ap; a.w = u_ext + L0.lo; a.fin = FIN_STORE; a.out = D->partial_send; a.gated = true;
    if (!(pre && D->first_replicated == 0)) wait_for_ghosts(c, a);
    if (pre && D->first_replicated == 0) { a.x = (u_ext + L0.lo) - L0.off; a.x_lo = L0.off; a.x_hi = L0.off + L0.own; }
    else set_shifted_x(a, u_ext, L0.off, L0.lo, L0.own, L0.hi);
    MGCG_CHECK(spmv_launch(c, c->planA, EPI_DOT, a));
    dist_trace(c, "A u + u.Au");
    MGCG_CHECK(dist_reduce(c, pre ? DF_CGCG : DF_CGCG_PLAIN, 3, true));
    dist_trace(c, "reduce w.u, r.u, r.r");
    return MGCG_OK;
}

// phase -> total device time between consecutive trace events (the label names the phase that ENDS at the event)
static void dist_trace_dump(mgcg_ctx *c, int iterations, float solve_ms)
{
    DistState *D = c->dist;
    std::vector<std::pair<const char *, cudaEvent_t>> &ev = D->trace_ev;
    std::vector<std::pair<std::string, std::pair<double, int>>> acc;
    for (size_t i = 1; i < ev.size(); ++i) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, ev[i - 1].second, ev[i].second) != cudaSuccess) { cudaGetLastError(); continue; }
        size_t j = 0;
        for (; j < acc.size(); ++j) if (acc[j].first == ev[i].first) break;
        if (j == acc.size()) acc.push_back({ev[i].first, {0.0, 0}});
        acc[j].second.first += t; acc[j].second.second++;
    }
    if (D->rank == 0) {
        fprintf(stderr, "[mgcg dist trace] rank 0 of %d, %d iterations, solve %.3f ms (eager launches; includes the speculative tail)\n", D->world, iterations, solve_ms);
        for (auto &a : acc)
            fprintf(stderr, "  %-34s n=%5d total %9.3f ms  avg %8.2f us\n", a.first.c_str(), a.second.second, a.second.first, 1e3 * a.second.first / a.second.second);
    }
    for (auto &e : ev) cudaEventDestroy(e.second);
    ev.clear();
}

static int dist_enqueue_thunk(mgcg_ctx *c, void *x)
{
    if (c->opt.cg_variant == MGCG_CG_SINGLE_REDUCE) return dist_enqueue_iteration_cgcg(c, (double *)x);
    return dist_enqueue_iteration(c, (double *)x);
}

static int dist_solve(mgcg_ctx *c, const double *b, double *x, mgcg_result *res)
{
    DistState *D = c->dist;
    if (!D || D->L.empty()) { set_error("mgcg_solve_dist: mgcg_setup_dist has not been called"); return MGCG_ERR_INVALID; }
    if (!c->is_setup) {
        set_error("mgcg_solve_dist: no valid set-up (the last mgcg_setup_dist failed, or a solve ended in MGCG_ERR_COMM and left the "
                  "exchange counters of the ranks out of step): call mgcg_setup_dist again on every rank");
        return MGCG_ERR_INVALID;
    }
    DistLevel &L0 = D->L[0];
    const int n = L0.own;
    const bool pre = (c->opt.precond == MGCG_PRECOND_MG);
    cudaStream_t st = c->stream;
    struct Events {   // destroyed on every exit path
        cudaEvent_t e[4] = {nullptr, nullptr, nullptr, nullptr};
        ~Events() { for (auto &x : e) if (x) cudaEventDestroy(x); }
    } evs;
    MGCG_CUDA(cudaEventCreate(&evs.e[0])); MGCG_CUDA(cudaEventCreate(&evs.e[1]));
    MGCG_CUDA(cudaEventCreateWithFlags(&evs.e[2], cudaEventDisableTiming));
    MGCG_CUDA(cudaEventCreateWithFlags(&evs.e[3], cudaEventDisableTiming));
    cudaEvent_t &ev0 = evs.e[0], &ev1 = evs.e[1];
    cudaEvent_t *evp = &evs.e[2];
    c->launches = 0; D->halo_exchanges = 0; D->allgathers = 0;
    if (D->coarse) D->coarse->launches = 0;
    Scalars h{};
    h.tolerance = c->opt.tolerance; h.min_iteration = c->opt.min_iteration; h.max_iteration = effective_max_iteration(c->opt.max_iteration, D->n_glob);
    h.norm = c->opt.norm; h.min_strict = c->opt.min_strict;
    c->h_scalars[0] = h;
    MGCG_CUDA(cudaMemcpyAsync(c->d_scalars, &c->h_scalars[0], sizeof(Scalars), cudaMemcpyHostToDevice, st));
    MGCG_CUDA(cudaEventRecord(ev0, st));
    double *p_own = D->p + L0.lo, *r_own = D->r + L0.lo;
    // rendezvous: ranks enter a solve with arbitrary host skew; the first wait of the peer transport must not time a late rank out
    if (D->peer.active) MGCG_CHECK(peer_reduce(c, -1, 1, false, 64));
    // init: Ap = A x0 (x0 staged through p's ghosts), r = b - Ap, p = r or M r, r.r [, r.z]   (Solve0, cs:436-463)
    MGCG_CUDA(cudaMemcpyAsync(p_own, x, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, st));
    MGCG_CHECK(dist_halo_exchange(c, D->p, n, L0.lo, L0.hi, L0.halo));
    {
        SpmvArgs a;
        a.A = c->A; set_shifted_x(a, D->p, L0.off, L0.lo, L0.own, L0.hi); a.y = D->Ap;
        MGCG_CHECK(spmv_launch(c, c->planA, EPI_PLAIN, a));
    }
    const bool cgcg = c->opt.cg_variant == MGCG_CG_SINGLE_REDUCE;
    if (!pre) {
        MGCG_CHECK(cg_init_residual(c, n, b, D->Ap, r_own, cgcg ? nullptr : p_own, FIN_STORE, D->partial_send));
        MGCG_CHECK(dist_reduce(c, DF_INIT_PLAIN, 1, false));
    } else {
        MGCG_CHECK(cg_init_residual(c, n, b, D->Ap, r_own, nullptr, FIN_STORE, D->partial_send));
        double *z = nullptr;
        MGCG_CHECK(dist_vcycle(c, D->r, &z, false, D->partial_send + 1));
        MGCG_CHECK(dist_reduce(c, DF_INIT_PRE, 2, false));
        if (cgcg) D->cgcg_u = z;
        else MGCG_CHECK(vec_copy(c, n, z, p_own));
    }
    // fused halo exchange: inside the loop p's ghost rows are pushed by the kernel that updates p; the first iteration's come from here
    if (!cgcg && dist_fused(c)) MGCG_CHECK(dist_halo_exchange(c, D->p, n, L0.lo, L0.hi, L0.halo));
    if (cgcg) {   // w = A u ; delta = w.u ; alpha = gamma / delta ; beta = 0 ; p = s = 0
        if (!D->s) MGCG_CUDA(pool_alloc((void **)&D->s, sizeof(double) * ((size_t)n + 1), st));
        const bool repl = pre && D->first_replicated == 0;
        double *u_ext = pre ? D->cgcg_u - L0.lo : D->r;
        if (!repl) MGCG_CHECK(dist_halo_exchange(c, u_ext, n, L0.lo, L0.hi, L0.halo));
        SpmvArgs a;
        a.A = c->A; a.y = D->Ap; a.w = u_ext + L0.lo; a.fin = FIN_STORE; a.out = D->partial_send;
        if (repl) { a.x = (u_ext + L0.lo) - L0.off; a.x_lo = L0.off; a.x_hi = L0.off + L0.own; }
        else set_shifted_x(a, u_ext, L0.off, L0.lo, L0.own, L0.hi);
        MGCG_CHECK(spmv_launch(c, c->planA, EPI_DOT, a));
        MGCG_CHECK(dist_reduce(c, DF_CGCG_INIT, 1, false));
        MGCG_CUDA(cudaMemsetAsync(p_own, 0, sizeof(double) * (size_t)n, st));
        MGCG_CUDA(cudaMemsetAsync(D->s, 0, sizeof(double) * (size_t)n, st));
    }
    const int interval = c->opt.check_interval > 0 ? c->opt.check_interval : 8;
    long long enq = 0;
    int slot = 0;
    bool finished = false, have_prev = false;
    const long long hard_cap = (long long)h.max_iteration + 2;
    // peer transport: the iteration is made of this library's kernels only, so it is captured once and replayed
    {
        const char *e = getenv("MGCG_DIST_TRACE");
        D->trace = e && e[0] == '1';
    }
    const bool use_graph = D->peer.active && c->opt.use_graph && !c->opt.profile && !D->trace;
    const int want_mode = use_graph ? (c->opt.use_graph >= 2 ? 2 : 1) : 0;
    if (use_graph && (!D->graph_exec || D->graph_b != b || D->graph_x != x || D->graph_variant != c->opt.cg_variant ||
                      (D->graph_mode != want_mode && D->graph_mode != -want_mode))) {
        dist_graph_free(D);
        D->graph_variant = c->opt.cg_variant;
        const int l0 = c->launches, lc0 = D->coarse ? D->coarse->launches : 0, h0 = D->halo_exchanges, g0 = D->allgathers;
        D->graph_mode = 0;
        if (want_mode == 2) {   // WHILE(!done) { iteration } on the device: every rank leaves the loop at the same iteration
            int s = graph_build_while(c, dist_enqueue_thunk, x, &D->graph, &D->graph_exec, &D->graph_nodes);
            if (s == MGCG_OK) D->graph_mode = 2;
            else if (s != MGCG_ERR_UNSUPPORTED) return s;
        }
        if (D->graph_mode == 0) {
            c->launches = l0; D->halo_exchanges = h0; D->allgathers = g0;
            if (D->coarse) D->coarse->launches = lc0;
            MGCG_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            int s = dist_enqueue_thunk(c, x);
            cudaError_t e = cudaStreamEndCapture(st, &D->graph);
            if (s != MGCG_OK) return s;
            MGCG_CUDA(e);
            MGCG_CUDA(cudaGraphInstantiate(&D->graph_exec, D->graph, 0));
            size_t nodes = 0;
            cudaGraphGetNodes(D->graph, nullptr, &nodes);
            D->graph_nodes = (int)nodes;
            D->graph_mode = want_mode == 2 ? -2 : 1;
        }
        D->graph_halo = D->halo_exchanges - h0; D->graph_gather = D->allgathers - g0;
        D->graph_b = b; D->graph_x = x;
        c->launches = l0; D->halo_exchanges = h0; D->allgathers = g0;
        if (D->coarse) D->coarse->launches = lc0;
    }
    const bool while_loop = use_graph && D->graph_mode == 2;
    if (while_loop) {
        MGCG_CUDA(cudaGraphLaunch(D->graph_exec, st));
        finished = true;
    }
    while (!finished) {   // every rank sees identical scalars, so every rank enqueues the same exchanges
        for (int k = 0; k < interval; ++k) {
            if (use_graph) {
                MGCG_CUDA(cudaGraphLaunch(D->graph_exec, st));
                c->launches += D->graph_nodes; D->halo_exchanges += D->graph_halo; D->allgathers += D->graph_gather;
            } else {
                MGCG_CHECK(dist_enqueue_thunk(c, x));
            }
            ++enq;
        }
        MGCG_CUDA(cudaMemcpyAsync(&c->h_scalars[slot], c->d_scalars, sizeof(Scalars), cudaMemcpyDeviceToHost, st));
        MGCG_CUDA(cudaEventRecord(evp[slot], st));
        if (have_prev) {
            MGCG_CUDA(cudaEventSynchronize(evp[slot ^ 1]));
            if (c->h_scalars[slot ^ 1].done) finished = true;
        }
        have_prev = true;
        slot ^= 1;
        if (!finished && enq > hard_cap + 2LL * interval) finished = true;
    }
    MGCG_CUDA(cudaMemcpyAsync(&c->h_scalars[0], c->d_scalars, sizeof(Scalars), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaEventRecord(ev1, st));
    MGCG_CUDA(cudaEventSynchronize(ev1));
    const Scalars fin = c->h_scalars[0];
    if (while_loop) {
        const int its = fin.iteration + 1;
        c->launches += D->graph_nodes * its; D->halo_exchanges += D->graph_halo * its; D->allgathers += D->graph_gather * its;
    }
    float ms = 0.f;
    MGCG_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));
    if (res) {
        memset(res, 0, sizeof(*res));
        res->status = fin.done ? fin.status : MGCG_ERR_NOT_CONVERGED;
        res->iteration = fin.iteration; res->iterations_done = fin.iteration + 1;
        res->residual = fin.residual; res->rr = fin.rr; res->rr0 = fin.rr0;
        res->solve_ms = ms; res->setup_ms = c->setup_ms;
        res->levels = pre ? D->first_replicated + mgcg_mg_levels(D->coarse) : 1;
        res->kernel_launches = c->launches + (D->coarse ? D->coarse->launches : 0);
        res->operator_complexity = 1.0;
        res->reserved[0] = D->halo_exchanges; res->reserved[1] = D->allgathers; res->reserved[2] = D->first_replicated;
        res->reserved[3] = D->peer.active ? MGCG_TRANSPORT_PEER : MGCG_TRANSPORT_NCCL; res->reserved[4] = use_graph ? (while_loop ? 2 : 1) : 0;
    }
    if (D->trace) dist_trace_dump(c, fin.iteration + 1, ms);
    if (!fin.done) { set_error("mgcg_solve_dist: iteration cap reached without the device flag"); return MGCG_ERR_NOT_CONVERGED; }
    if (fin.status != MGCG_OK)
        set_error(fin.status == MGCG_ERR_NOT_CONVERGED ? "mgcg_solve_dist: Iteration > MaxIteration"
                  : fin.status == MGCG_ERR_COMM ? "mgcg_solve_dist: a peer-memory exchange timed out inside the loop (a rank fell behind or died); the set-up is void"
                  : "mgcg_solve_dist: breakdown (p.Ap <= 0 or NaN)");
    return fin.status;
}

}  // namespace mgcg

using namespace mgcg;

extern "C" int mgcg_comm_unique_id(void *out128)
{
    if (!out128) { set_error("mgcg_comm_unique_id: null argument"); return MGCG_ERR_INVALID; }
    NcclApi *a = nullptr;
    MGCG_CHECK(nccl_load(&a));
    ncclUniqueId id;
    MGCG_NCCL(a, GetUniqueId(&id));
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    memcpy(out128, &id, 128);
    return MGCG_OK;
}

extern "C" int mgcg_comm_init(mgcg_ctx *c, const void *id128, int rank, int world)
{
    if (!c || !id128 || world < 1 || world > kMaxRanks || rank < 0 || rank >= world) {
        set_error("mgcg_comm_init: bad argument (world must be 1..%d)", kMaxRanks);
        return MGCG_ERR_INVALID;
    }
    MGCG_CUDA(cudaSetDevice(c->device));
    dist_free(c);
    DistState *D = new DistState();
    c->dist = D;
    MGCG_CHECK(nccl_load(&D->api));
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    D->rank = rank; D->world = world;
    // NCCL carries the setup's exchanges only (a few MB of ghost rows and the replicated level); the loop runs on the
    // library's own peer-memory kernels.  So the communicator is asked for few channels: connecting them is what the first
    // collective / first send to a peer pays for, per channel and per peer.  MGCG_NCCL_MAX_CTAS=0 keeps NCCL's default.
    int max_ctas = 4;
    if (const char *e = getenv("MGCG_NCCL_MAX_CTAS")) max_ctas = atoi(e);
    if (D->api->CommInitRankConfig && max_ctas > 0) {
        ncclConfig_t cfg = NCCL_CONFIG_INITIALIZER;
        cfg.minCTAs = 1;
        cfg.maxCTAs = max_ctas;
        MGCG_NCCL(D->api, CommInitRankConfig(&D->comm, world, id, rank, &cfg));
    } else {
        MGCG_NCCL(D->api, CommInitRank(&D->comm, world, id, rank));
    }
    MGCG_CUDA(cudaMalloc((void **)&D->partial_send, 4 * sizeof(double)));
    MGCG_CUDA(cudaMalloc((void **)&D->partial_all, 4 * sizeof(double) * (size_t)world));
    MGCG_CUDA(cudaMalloc((void **)&D->gather_ints, sizeof(int) * (size_t)(2 * world + 2)));
    MGCG_CUDA(cudaMemset(D->gather_ints, 0, sizeof(int) * (size_t)(2 * world + 2)));
    // open both connection classes now (once per process) instead of inside the first setup
    if (world > 1 && !getenv("MGCG_NCCL_LAZY")) {
        int *g = D->gather_ints;
        MGCG_NCCL(D->api, AllGather(g + world, g, 1, ncclInt, D->comm, c->stream));
        MGCG_NCCL(D->api, GroupStart());
        for (int q = 0; q < world; ++q) {
            if (q == rank) continue;
            MGCG_NCCL(D->api, Send(g + world, 1, ncclInt, q, D->comm, c->stream));
            MGCG_NCCL(D->api, Recv(g + world + 1 + q, 1, ncclInt, q, D->comm, c->stream));
        }
        MGCG_NCCL(D->api, GroupEnd());
        MGCG_CUDA(cudaStreamSynchronize(c->stream));
    }
    return MGCG_OK;
}

extern "C" int mgcg_setup_dist(mgcg_ctx *c, int global_count, int row_offset, int local_count, int local_elements_count,
                               const double *d_elements, const int *d_row_offsets, const int *d_column_indeces,
                               const mgcg_options *opt)
{
    if (!c || !c->dist) { set_error("mgcg_setup_dist: call mgcg_comm_init first"); return MGCG_ERR_INVALID; }
    if (!d_elements || !d_row_offsets || !d_column_indeces || local_count <= 0 || global_count <= 0) {
        set_error("mgcg_setup_dist: null or empty argument (every rank must own at least one row)");
        return MGCG_ERR_INVALID;
    }
    MGCG_CUDA(cudaSetDevice(c->device));
    if (opt) c->opt = *opt;
    if (c->opt.norm == MGCG_NORM_ABS_LINF) { set_error("mgcg_setup_dist: the max-norm stop rule is not available on the sharded path"); return MGCG_ERR_UNSUPPORTED; }
    if (c->opt.mg_smoother != MGCG_SMOOTHER_JACOBI && c->opt.mg_smoother != MGCG_SMOOTHER_CHEBYSHEV) {
        set_error("mgcg_setup_dist: the sharded levels are smoothed by weighted Jacobi or Chebyshev (red-black Gauss-Seidel is single-GPU)"); return MGCG_ERR_UNSUPPORTED;
    }
    if (c->opt.precond == MGCG_PRECOND_MG && c->opt.mg_aggregation == MGCG_AGG_MIS2) {
        set_error("mgcg_setup_dist: MIS-2 aggregation is single-GPU only (sharded levels aggregate by grid blocks or blocks of 8 rows)");
        return MGCG_ERR_UNSUPPORTED;
    }
    c->is_setup = false;
    cudaEvent_t e0, e1;
    MGCG_CUDA(cudaEventCreate(&e0)); MGCG_CUDA(cudaEventCreate(&e1));
    MGCG_CUDA(cudaEventRecord(e0, c->stream));
    int s = dist_setup(c, global_count, row_offset, local_count, local_elements_count, d_elements, d_row_offsets, d_column_indeces);
    if (s == MGCG_OK) {
        cudaEventRecord(e1, c->stream);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&c->setup_ms, e0, e1);
        c->is_setup = true;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return s;
}

extern "C" int mgcg_solve_dist(mgcg_ctx *c, const double *d_b_local, double *d_x_local, mgcg_result *res)
{
    if (!c || !d_b_local || !d_x_local) { set_error("mgcg_solve_dist: null argument"); return MGCG_ERR_INVALID; }
    MGCG_CUDA(cudaSetDevice(c->device));
    int s = dist_solve(c, d_b_local, d_x_local, res);
    // A wait that timed out inside the loop (halo.cuh halo_wait) ends the solve with blocks that never reached halo_signal: the
    // ticket of the exchange and this rank's sequence number no longer agree with the peers'.  Only peer_open resets them, so the
    // set-up is void until mgcg_setup_dist runs again — the next solve must not read ghost rows through stale counters.
    if (s == MGCG_ERR_COMM) c->is_setup = false;
    if (res) res->status = s;
    return s;
}

// ---- sharded solve on HOST pointers: this rank's row block in host arrays (the reference's multi-device class hands host
// arrays to the native side, ConjugateGradientParallelGpu.cs:358-379 -> Initialize, Mgcg.cu:57-85: elements / column indices
// of the block, row offsets of the block — rebased here like Mgcg.cu:73 when they are not — x and b slices), one call:
// upload + hierarchy build + solve + download.  Device copies are owned by the context and reused by the next call that fits.
namespace mgcg {
struct HostPathDist {
    double *val = nullptr; int *rowptr = nullptr; int *col = nullptr; double *b = nullptr; double *x = nullptr;
    size_t cap_n = 0, cap_nnz = 0;
    void release()
    {
        cudaFree(val); cudaFree(rowptr); cudaFree(col); cudaFree(b); cudaFree(x);
        val = nullptr; rowptr = nullptr; col = nullptr; b = nullptr; x = nullptr; cap_n = cap_nnz = 0;
    }
};
__global__ void k_rebase_rowptr(int n, int *rp, int base)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) rp[i] -= base;
}
bool dist_x_range(const mgcg_ctx *c, int *x_lo, int *x_hi)
{
    const DistState *D = c->dist;
    if (!D || D->L.empty()) return false;
    const DistLevel &L0 = D->L[0];
    *x_lo = L0.off - L0.lo; *x_hi = L0.off + L0.own + L0.hi;
    return true;
}

void dist_host_path_free(mgcg_ctx *c)
{
    if (!c->host_path_dist) return;
    HostPathDist *h = (HostPathDist *)c->host_path_dist;
    h->release();
    delete h;
    c->host_path_dist = nullptr;
}
}  // namespace mgcg

extern "C" int mgcg_solve_host_dist(mgcg_ctx *c, int global_count, int row_offset, int local_count, int local_elements_count,
                                    const double *elements, const int *row_offsets, const int *column_indeces,
                                    const double *b_local, double *x_local, const mgcg_options *opt, mgcg_result *res)
{
    if (!c || !c->dist) { set_error("mgcg_solve_host_dist: call mgcg_comm_init first"); return MGCG_ERR_INVALID; }
    if (!elements || !row_offsets || !column_indeces || !b_local || !x_local || local_count <= 0 || local_elements_count < 0) {
        set_error("mgcg_solve_host_dist: null or empty argument");
        return MGCG_ERR_INVALID;
    }
    MGCG_CUDA(cudaSetDevice(c->device));
    if (!c->host_path_dist) c->host_path_dist = new HostPathDist();
    HostPathDist *h = (HostPathDist *)c->host_path_dist;
    const size_t n = (size_t)local_count, nnz = (size_t)local_elements_count;
    if (h->cap_n < n || h->cap_nnz < nnz) {
        h->release();
        MGCG_CUDA(cudaMalloc((void **)&h->val, sizeof(double) * (nnz > 0 ? nnz : 1)));
        MGCG_CUDA(cudaMalloc((void **)&h->col, sizeof(int) * (nnz > 0 ? nnz : 1)));
        MGCG_CUDA(cudaMalloc((void **)&h->rowptr, sizeof(int) * (n + 1)));
        MGCG_CUDA(cudaMalloc((void **)&h->b, sizeof(double) * n));
        MGCG_CUDA(cudaMalloc((void **)&h->x, sizeof(double) * n));
        h->cap_n = n; h->cap_nnz = nnz;
    }
    cudaStream_t st = c->stream;
    MGCG_CUDA(cudaMemcpyAsync(h->val, elements, sizeof(double) * nnz, cudaMemcpyHostToDevice, st));
    MGCG_CUDA(cudaMemcpyAsync(h->col, column_indeces, sizeof(int) * nnz, cudaMemcpyHostToDevice, st));
    MGCG_CUDA(cudaMemcpyAsync(h->rowptr, row_offsets, sizeof(int) * (n + 1), cudaMemcpyHostToDevice, st));
    if (row_offsets[0] != 0) {   // offsets of the block inside the global array: rebase to the block (Mgcg.cu:73)
        k_rebase_rowptr<<<(local_count + 256) / 256, 256, 0, st>>>(local_count + 1, h->rowptr, row_offsets[0]);
        MGCG_LAUNCH_CHECK();
    }
    MGCG_CUDA(cudaMemcpyAsync(h->b, b_local, sizeof(double) * n, cudaMemcpyHostToDevice, st));
    MGCG_CUDA(cudaMemcpyAsync(h->x, x_local, sizeof(double) * n, cudaMemcpyHostToDevice, st));
    MGCG_CHECK(mgcg_setup_dist(c, global_count, row_offset, local_count, local_elements_count, h->val, h->rowptr, h->col, opt));
    const int s = mgcg_solve_dist(c, h->b, h->x, res);
    MGCG_CUDA(cudaMemcpyAsync(x_local, h->x, sizeof(double) * n, cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    return s;
}

// Collective teardown of the peer transport: every rank unmaps its peers' arenas, all ranks rendezvous, then each frees its own
// arena (cudaFree of memory another process still has open is undefined).  Call on every rank before mgcg_destroy
// when the ranks are separate processes; mgcg_destroy alone frees without a rendezvous.
extern "C" int mgcg_comm_finalize(mgcg_ctx *c)
{
    if (!c) { set_error("mgcg_comm_finalize: null context"); return MGCG_ERR_INVALID; }
    if (!c->dist) return MGCG_OK;
    MGCG_CUDA(cudaSetDevice(c->device));
    dist_graph_free(c->dist);
    peer_close(c, true);
    c->is_setup = false;
    return MGCG_OK;
}
