// peer.cu — peer-memory transport of the sharded solve: halo exchange, coarse right-hand-side all-gather and the scalar
// all-reductions done by this library's own kernels with stores into the neighbours' HBM over NVLink / NVSwitch.
//
// Replaces, inside the CG loop, the grouped ncclSend/ncclRecv + ncclAllGather calls of the first sharded version (which
// themselves replaced SyncP = P2Host + P2Device through a pinned host array, ConjugateGradientParallelGpu.cs:384-419,
// Mgcg.cu:88-113, and the host-side `resultsDot.Sum()`, cs:463,499,525).  At 256^3 on 8 GPUs an iteration is ~0.25 ms of
// sharded compute against ~11 NCCL operations of 15-25 us each, none of them capturable cheaply: the loop was latency
// bound.  Here an exchange is ONE kernel of ours (~3-5 us) and the whole iteration is captured into a CUDA graph.
//
// Protocol.  Every rank owns an ARENA (cudaMalloc, exported with cudaIpcGetMemHandle and mapped by every other rank;
// ranks that live in one process pass the raw pointer and enable peer access instead):
//     PeerCtl   flags and the slots of the scalar reductions
//     stage[2]  receive staging, double-buffered by the parity of the exchange sequence number
// An exchange `s` (same number on every rank, kept on the device so a captured graph can be replayed) is
//     phase 1  copy my send spans into the receivers' stage[s & 1] (plain stores to peer memory), __threadfence_system,
//              and the block that takes the last ticket publishes  receiver.xflag[my rank] = s  (st.release.sys)
//     phase 2  wait (ld.acquire.sys) until xflag[q] >= s for every rank q I receive from, then unpack stage[s & 1]
//              into the ghost rows of the destination vector.
// Double buffering is what makes it safe without an acknowledgement: a NEIGHBOUR can overwrite its halo slot of
// stage[s & 1] only in exchange s+2, which it reaches after waiting for MY flag of exchange s+1, which I publish after my
// unpack of s.  The all-gather is pushed by every rank (not only neighbours): it has its own staging region, and two
// gathers are always separated by a scalar reduction, which is a full barrier (every rank waits for every rank).
// The scalar reduction is the same idea in one block: every rank stores its partial sums into slot[s & 1][rank] of every
// rank, then sums the slots IN RANK ORDER (deterministic and identical everywhere, like the reference's host sum).
// A wait that lasts longer than ~4 s marks the solve as failed (MGCG_ERR_COMM) instead of hanging the device.
#include "dist.cuh"
#include "reduce.cuh"
#include <unistd.h>

namespace mgcg {

// Clocks a wait may last before the solve is marked MGCG_ERR_COMM instead of hanging the device: ~4 s at 2 GHz by default,
// MGCG_PEER_TIMEOUT_S overrides (peer_open).  The rendezvous at the start of every solve waits 64 x longer: ranks enter a
// solve with whatever skew their hosts have (a first-call graph instantiation, a profiler, a debugger), and a rank that is
// merely late must not fail its peers — inside the loop the ranks are already in step and a long wait means a dead peer.
__device__ long long g_peer_wait_clocks = 8000000000ll;

// true when the flag reached `seq`; false after the time limit
__device__ __forceinline__ bool wait_flag(const unsigned long long *flag, unsigned long long seq, long long scale = 1)
{
    const long long t0 = clock64();
    const long long limit = g_peer_wait_clocks * scale;
    while (ld_acquire_sys(flag) < seq) {
        if (clock64() - t0 > limit) return false;
        __nanosleep(32);
    }
    return true;
}

__device__ __forceinline__ void copy_span(double *__restrict__ dst, const double *__restrict__ src, int count, int gtid, int gsize,
                                          bool src_remote_written)
{
    if (count <= 0) return;
    if (((((uintptr_t)dst) | ((uintptr_t)src)) & 15u) == 0) {
        const int n2 = count >> 1;
        const double2 *s2 = reinterpret_cast<const double2 *>(src);
        double2 *d2 = reinterpret_cast<double2 *>(dst);
        // staging is written by another GPU: read it through L2 (never from a stale L1 line)
        for (int i = gtid; i < n2; i += gsize) d2[i] = src_remote_written ? __ldcg(s2 + i) : s2[i];
        if ((count & 1) && gtid == 0) dst[count - 1] = src_remote_written ? __ldcg(src + count - 1) : src[count - 1];
    } else {
        for (int i = gtid; i < count; i += gsize) dst[i] = src_remote_written ? __ldcg(src + i) : src[i];
    }
}

__global__ void __launch_bounds__(256) k_peer_exchange(XchgDesc d, PeerCtl *ctl, Scalars *sc, bool gated)
{
    if (gated && sc->done) return;
    __shared__ int s_last;
    __shared__ int s_fail;
    const int t = threadIdx.x;
    const int gtid = blockIdx.x * blockDim.x + t, gsize = gridDim.x * blockDim.x;
    const unsigned long long seq = ld_volatile_u64(&ctl->xseq) + 1ull;   // every block reads it before the last one advances it
    const size_t par = (size_t)(seq & 1ull) * d.stage_stride;
    if (t == 0) s_fail = 0;
    // ---- phase 1: push
    for (int i = 0; i < d.ns; ++i) copy_span(d.s[i].dst + par, d.s[i].src, d.s[i].count, gtid, gsize, false);
    __threadfence_system();
    __syncthreads();
    if (t == 0) {
        const unsigned int k = atomicAdd(&ctl->ticket1, 1u);
        s_last = (k == gridDim.x - 1);
    }
    __syncthreads();
    if (s_last && t < d.ns) {
        __threadfence_system();
        st_release_sys(d.s[t].flag, seq);
    }
    // ---- phase 2: wait for the senders, unpack
    if (t < d.nr && !wait_flag(d.r[t].flag, seq)) s_fail = 1;
    __syncthreads();
    for (int i = 0; i < d.nr; ++i) copy_span(d.r[i].dst, d.r[i].stage + par, d.r[i].count, gtid, gsize, true);
    __syncthreads();
    if (t == 0) {
        if (s_fail) { sc->status = MGCG_ERR_COMM; sc->done = 1; }
        __threadfence();
        const unsigned int k = atomicAdd(&ctl->ticket2, 1u);
        if (k == gridDim.x - 1) {   // every block has read xseq and finished: advance
            ctl->ticket1 = 0u; ctl->ticket2 = 0u;
            __threadfence();
            ctl->xseq = seq;
        }
    }
}

// scalar all-reduction + the CG bookkeeping that follows it, one block.  `mine` holds this rank's k partial sums.
__global__ void __launch_bounds__(64) k_peer_reduce(int mode, int k, int rank, int world, PeerPtrs pp, const double *__restrict__ mine,
                                                   Scalars *sc, bool gated, int wait_scale)
{
    if (gated && sc->done) return;
    __shared__ double s_v[kPeerMax][3];
    __shared__ int s_fail;
    const int t = threadIdx.x;
    PeerCtl *me = pp.ctl[rank];
    const unsigned long long seq = ld_volatile_u64(&me->rseq) + 1ull;
    const int par = (int)(seq & 1ull);
    if (t == 0) s_fail = 0;
    __syncthreads();
    if (t < world) {
        PeerCtl *q = pp.ctl[t];
        q->rslot[par][rank][0] = mine[0];
        q->rslot[par][rank][1] = k > 1 ? mine[1] : 0.0;
        q->rslot[par][rank][2] = k > 2 ? mine[2] : 0.0;
        __threadfence_system();
        st_release_sys(&q->rflag[rank], seq);
        if (!wait_flag(&me->rflag[t], seq, wait_scale)) s_fail = 1;
        s_v[t][0] = __ldcg(&me->rslot[par][t][0]);
        s_v[t][1] = __ldcg(&me->rslot[par][t][1]);
        s_v[t][2] = __ldcg(&me->rslot[par][t][2]);
    }
    __syncthreads();
    if (t != 0) return;
    if (s_fail) { sc->status = MGCG_ERR_COMM; sc->done = 1; me->rseq = seq; return; }
    double s0 = 0.0, s1 = 0.0, s2 = 0.0;
    for (int r = 0; r < world; ++r) { s0 += s_v[r][0]; s1 += s_v[r][1]; s2 += s_v[r][2]; }   // rank order (cs:463 resultsDot.Sum())
    if (mode >= 0) dist_finalize(mode, s0, s1, s2, sc);
    me->rseq = seq;
}

// ---------------------------------------------------------------------------------------------------------
static int grid_for(long long doubles, int sm_count)
{
    static int per_block = 0;
    if (!per_block) { const char *e = getenv("MGCG_PEER_DOUBLES_PER_BLOCK"); per_block = e ? atoi(e) : 1024; if (per_block < 256) per_block = 256; }   // measured per 256^2-double halo: 1024 -> 13.6 us, 4096 -> 16.1, 16384 -> 27.8
    long long g = (doubles + per_block - 1) / per_block;   // >= 16 doubles (2 x 128-bit stores in flight... per thread at 256 threads)
    if (g < 1) g = 1;
    if (g > sm_count) g = sm_count;           // all blocks must be co-resident: they spin on the flags
    return (int)g;
}

int peer_exchange_launch(mgcg_ctx *c, const XchgDesc &d, bool gated)
{
    DistState *D = c->dist;
    long long vol = 0;
    for (int i = 0; i < d.ns; ++i) vol += d.s[i].count;
    long long rv = 0;
    for (int i = 0; i < d.nr; ++i) rv += d.r[i].count;
    if (rv > vol) vol = rv;
    k_peer_exchange<<<grid_for(vol, c->sm_count), 256, 0, c->stream>>>(d, (PeerCtl *)D->peer.arena, c->d_scalars, gated);
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

int peer_halo_exchange(mgcg_ctx *c, double *v, int own, int lo, int hi, int halo)
{
    DistState *D = c->dist;
    PeerState &P = D->peer;
    if (halo > P.halo_cap) { set_error("peer_halo_exchange: halo %d exceeds the staging capacity %d", halo, P.halo_cap); return MGCG_ERR_INVALID; }
    XchgDesc d;
    memset(&d, 0, sizeof(d));
    d.stage_stride = P.stage_cap;
    double *own0 = v + lo;
    const int rk = D->rank, W = D->world;
    auto stage0 = [&](int q) { return (double *)((char *)P.peer[q] + P.stage_off); };
    auto ctl = [&](int q) { return (PeerCtl *)P.peer[q]; };
    if (rk > 0) {            // my first `halo` rows are rank-1's high ghosts
        d.s[d.ns].src = own0; d.s[d.ns].dst = stage0(rk - 1) + P.halo_cap; d.s[d.ns].flag = &ctl(rk - 1)->xflag[rk]; d.s[d.ns].count = halo; d.ns++;
        d.r[d.nr].stage = stage0(rk); d.r[d.nr].dst = v; d.r[d.nr].flag = &ctl(rk)->xflag[rk - 1]; d.r[d.nr].count = lo; d.nr++;
    }
    if (rk < W - 1) {
        d.s[d.ns].src = own0 + own - halo; d.s[d.ns].dst = stage0(rk + 1); d.s[d.ns].flag = &ctl(rk + 1)->xflag[rk]; d.s[d.ns].count = halo; d.ns++;
        d.r[d.nr].stage = stage0(rk) + P.halo_cap; d.r[d.nr].dst = own0 + own; d.r[d.nr].flag = &ctl(rk)->xflag[rk + 1]; d.r[d.nr].count = hi; d.nr++;
    }
    return peer_exchange_launch(c, d, true);
}

// all-gather of a vector laid out by `offs` (own block already in place): every rank pushes its block to all others
int peer_allgather(mgcg_ctx *c, double *full, const int *offs)
{
    DistState *D = c->dist;
    PeerState &P = D->peer;
    const int rk = D->rank, W = D->world;
    if ((size_t)2 * P.halo_cap + (size_t)offs[W] > P.stage_cap) { set_error("peer_allgather: %d rows exceed the staging capacity", offs[W]); return MGCG_ERR_INVALID; }
    XchgDesc d;
    memset(&d, 0, sizeof(d));
    d.stage_stride = P.stage_cap;
    for (int q = 0; q < W; ++q) {
        if (q == rk) continue;
        double *stq = (double *)((char *)P.peer[q] + P.stage_off) + 2 * (size_t)P.halo_cap;
        double *stme = (double *)((char *)P.peer[rk] + P.stage_off) + 2 * (size_t)P.halo_cap;
        d.s[d.ns].src = full + offs[rk]; d.s[d.ns].dst = stq + offs[rk]; d.s[d.ns].flag = &((PeerCtl *)P.peer[q])->xflag[rk];
        d.s[d.ns].count = offs[rk + 1] - offs[rk]; d.ns++;
        d.r[d.nr].stage = stme + offs[q]; d.r[d.nr].dst = full + offs[q]; d.r[d.nr].flag = &((PeerCtl *)P.peer[rk])->xflag[q];
        d.r[d.nr].count = offs[q + 1] - offs[q]; d.nr++;
    }
    return peer_exchange_launch(c, d, true);
}

int peer_reduce(mgcg_ctx *c, int mode, int k, bool gated, int wait_scale)
{
    DistState *D = c->dist;
    PeerPtrs pp;
    for (int q = 0; q < kPeerMax; ++q) pp.ctl[q] = (PeerCtl *)(q < D->world ? D->peer.peer[q] : nullptr);
    k_peer_reduce<<<1, 64, 0, c->stream>>>(mode, k, D->rank, D->world, pp, D->partial_send, c->d_scalars, gated, wait_scale);
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
// arena life cycle
// Unmap the peers' arenas, then free this rank's own.  CUDA leaves freeing memory that another process still has open through
// cudaIpcOpenMemHandle undefined, and a peer may still be storing into it: `collective` = true (every rank is inside the same
// call: a re-setup, mgcg_comm_finalize) puts an NCCL all-reduce between the two steps, so that nobody frees before everybody
// has unmapped and drained.  Teardown through mgcg_destroy alone cannot assume the other ranks are there (one may have
// failed): callers that share arenas across processes call mgcg_comm_finalize on every rank first (include/mgcg.h).
void peer_close(mgcg_ctx *c, bool collective)
{
    DistState *D = c->dist;
    if (!D) return;
    PeerState &P = D->peer;
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (int q = 0; q < kPeerMax; ++q) {
        if (P.peer[q] && P.mapped[q]) cudaIpcCloseMemHandle(P.peer[q]);
        P.peer[q] = nullptr; P.mapped[q] = false;
    }
    if (collective && P.arena && D->comm && D->api && D->world > 1) {
        int *d_flag = nullptr;
        if (cudaMalloc((void **)&d_flag, sizeof(int)) == cudaSuccess) {
            cudaMemsetAsync(d_flag, 0, sizeof(int), c->stream);
            D->api->AllReduce(d_flag, d_flag, 1, ncclInt, ncclSum, D->comm, c->stream);
            cudaStreamSynchronize(c->stream);
            cudaFree(d_flag);
        }
    }
    if (P.arena) cudaFree(P.arena);
    P = PeerState();
}

struct PeerHello {
    cudaIpcMemHandle_t handle;
    long long pid;
    unsigned long long ptr;
    int device;
    int ok;
};

// Collective: allocate the arena for `halo_cap` ghost rows per side and `gather_cap` gathered rows, exchange the handles
// (over NCCL: this is setup, not the loop) and map every peer.  Leaves P.active = true only if EVERY rank succeeded.
int peer_open(mgcg_ctx *c, int halo_cap, int gather_cap, size_t vec_doubles)
{
    DistState *D = c->dist;
    PeerState &P = D->peer;
    cudaStream_t st = c->stream;
    const int W = D->world, rk = D->rank;
    P.active = false;
    if (W < 2 || W > kPeerMax) { peer_close(c, false); return MGCG_OK; }
    halo_cap = (halo_cap + 1) & ~1;
    // A re-setup keeps the arena and its mappings when the new hierarchy fits (allocating and IPC-mapping with 7 peers costs
    // tens of milliseconds): every rank clears its own flags / sequence numbers, and the all-reduce that takes the decision
    // is also the barrier that orders those memsets before anybody's first push.
    {
        int reuse = (P.ready && P.arena && halo_cap <= P.halo_cap && (size_t)2 * P.halo_cap + (size_t)gather_cap <= P.stage_cap &&
                     vec_doubles * sizeof(double) <= P.vec_cap) ? 1 : 0;
        int *d_flag = nullptr;
        MGCG_CUDA(pool_alloc((void **)&d_flag, sizeof(int), st));
        if (reuse) MGCG_CUDA(cudaMemsetAsync(P.arena, 0, P.stage_off, st));
        MGCG_CUDA(cudaMemcpyAsync(d_flag, &reuse, sizeof(int), cudaMemcpyHostToDevice, st));
        MGCG_NCCL(D->api, AllReduce(d_flag, d_flag, 1, ncclInt, ncclMin, D->comm, st));
        MGCG_CUDA(cudaMemcpyAsync(&reuse, d_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
        MGCG_CUDA(cudaStreamSynchronize(st));
        pool_release(d_flag);
        if (reuse) { P.active = true; P.fused = vec_doubles > 0; P.nvec = 0; P.vec_used = 0; return MGCG_OK; }
    }
    peer_close(c, true);      // collective: every rank takes this branch together (the all-reduce above decided it)
    {   // wait limit of the flag spins (seconds -> clocks at the device's clock rate)
        long long clocks = 8000000000ll;
        if (const char *e = getenv("MGCG_PEER_TIMEOUT_S")) {
            const double sec = atof(e);
            int khz = 0;
            cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, c->device);
            if (sec > 0 && khz > 0) clocks = (long long)(sec * 1e3 * (double)khz);
        }
        MGCG_CUDA(cudaMemcpyToSymbol(g_peer_wait_clocks, &clocks, sizeof(clocks)));
        P.wait_clocks = clocks;
    }
    // halo and gather staging are DISJOINT: a gather is pushed by every rank, also by ranks that are not this rank's
    // neighbours and have therefore not synchronised with it in the halo exchanges just before — their blocks must not land
    // where a slower rank may still be unpacking a halo of two exchanges ago (seen as a changed iteration count at 8 GPUs)
    size_t cap = (size_t)2 * halo_cap + (size_t)gather_cap;
    cap = (cap + 1) & ~(size_t)1;
    P.halo_cap = halo_cap; P.stage_cap = cap;
    P.stage_off = (sizeof(PeerCtl) + 255) & ~(size_t)255;
    P.vec_off = (P.stage_off + sizeof(double) * 2 * cap + 255) & ~(size_t)255;
    P.vec_cap = vec_doubles * sizeof(double);
    P.fused = vec_doubles > 0; P.nvec = 0; P.vec_used = 0;
    const size_t bytes = P.vec_off + P.vec_cap;
    PeerHello me;
    memset(&me, 0, sizeof(me));
    me.ok = 1;
    if (cudaMalloc(&P.arena, bytes) != cudaSuccess) { cudaGetLastError(); P.arena = nullptr; me.ok = 0; }
    if (me.ok) {
        MGCG_CUDA(cudaMemsetAsync(P.arena, 0, bytes, st));
        if (cudaIpcGetMemHandle(&me.handle, P.arena) != cudaSuccess) { cudaGetLastError(); me.ok = 0; }
    }
    me.pid = (long long)getpid(); me.ptr = (unsigned long long)(uintptr_t)P.arena; me.device = c->device;
    // hellos of every rank
    PeerHello *d_all = nullptr;
    MGCG_CUDA(cudaMalloc((void **)&d_all, sizeof(PeerHello) * (size_t)(W + 1)));
    MGCG_CUDA(cudaMemcpyAsync(d_all + W, &me, sizeof(me), cudaMemcpyHostToDevice, st));
    MGCG_NCCL(D->api, AllGather(d_all + W, d_all, sizeof(PeerHello), ncclChar, D->comm, st));
    std::vector<PeerHello> all((size_t)W);
    MGCG_CUDA(cudaMemcpyAsync(all.data(), d_all, sizeof(PeerHello) * (size_t)W, cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    int ok = 1;
    for (int q = 0; q < W; ++q) ok = ok && all[q].ok;
    if (ok) {
        for (int q = 0; q < W && ok; ++q) {
            if (q == rk) { P.peer[q] = P.arena; continue; }
            if (all[q].pid == me.pid) {   // same process (one host thread per GPU): the pointer itself, with peer access on
                int can = 0;
                if (all[q].device == c->device) can = 1;
                else if (cudaDeviceCanAccessPeer(&can, c->device, all[q].device) == cudaSuccess && can) {
                    cudaError_t e = cudaDeviceEnablePeerAccess(all[q].device, 0);
                    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) can = 0;
                    cudaGetLastError();
                }
                if (can) P.peer[q] = (void *)(uintptr_t)all[q].ptr; else ok = 0;
            } else {
                void *p = nullptr;
                if (cudaIpcOpenMemHandle(&p, all[q].handle, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess) { P.peer[q] = p; P.mapped[q] = true; }
                else { cudaGetLastError(); ok = 0; }
            }
        }
    }
    // agreement: everyone or no one.  The all-reduce is also the barrier that orders the arena memsets before any push.
    int *d_ok = (int *)d_all;
    MGCG_CUDA(cudaMemcpyAsync(d_ok, &ok, sizeof(int), cudaMemcpyHostToDevice, st));
    MGCG_NCCL(D->api, AllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, D->comm, st));
    MGCG_CUDA(cudaMemcpyAsync(&ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaStreamSynchronize(st));
    cudaFree(d_all);
    if (!ok) { peer_close(c, true); return MGCG_OK; }      // every rank sees the same `ok`
    P.active = true; P.ready = true;
    return MGCG_OK;
}

// ---------------------------------------------------------------------------------------------------------
// sharded vectors inside the arena (fused halo exchange, halo.cuh)
static inline size_t round256(size_t b) { return (b + 255) & ~(size_t)255; }

double *peer_vec_alloc(mgcg_ctx *c, int lo, int own, int hi, int halo)
{
    PeerState &P = c->dist->peer;
    if (!P.active || !P.fused || P.nvec >= kPeerVecMax) return nullptr;
    const size_t bytes = round256(sizeof(double) * ((size_t)lo + own + hi + 2));
    if (P.vec_used + bytes > P.vec_cap) return nullptr;
    double *v = (double *)((char *)P.arena + P.vec_off + P.vec_used);
    if (cudaMemsetAsync(v, 0, bytes, c->stream) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    const int i = P.nvec++;
    P.vbase[i] = v; P.vlo[i] = lo; P.vown[i] = own; P.vhi[i] = hi; P.vhalo[i] = halo;
    P.vec_used += bytes;
    return v;
}

struct PeerVecTable { int nvec, pad; PeerVecSlot slot[kPeerVecMax]; };

int peer_vec_publish(mgcg_ctx *c)
{
    DistState *D = c->dist;
    PeerState &P = D->peer;
    cudaStream_t st = c->stream;
    const int W = D->world, rk = D->rank;
    PeerVecTable me;
    memset(&me, 0, sizeof(me));
    me.nvec = P.nvec;
    for (int i = 0; i < P.nvec; ++i) {
        me.slot[i].lo_off = (size_t)((char *)P.vbase[i] - (char *)P.arena);
        me.slot[i].hi_off = me.slot[i].lo_off + sizeof(double) * ((size_t)P.vlo[i] + P.vown[i]);
    }
    PeerVecTable *d_all = nullptr;
    MGCG_CUDA(pool_alloc((void **)&d_all, sizeof(PeerVecTable) * (size_t)(W + 1), st));
    MGCG_CUDA(cudaMemcpyAsync(d_all + W, &me, sizeof(me), cudaMemcpyHostToDevice, st));
    MGCG_NCCL(D->api, AllGather(d_all + W, d_all, sizeof(PeerVecTable), ncclChar, D->comm, st));
    std::vector<PeerVecTable> all((size_t)W);
    MGCG_CUDA(cudaMemcpyAsync(all.data(), d_all, sizeof(PeerVecTable) * (size_t)W, cudaMemcpyDeviceToHost, st));
    MGCG_CUDA(cudaStreamSynchronize(st));     // (also: every rank's vectors are zeroed before anybody's first push)
    pool_release(d_all);
    for (int q = 0; q < W; ++q)
        if (all[q].nvec != P.nvec) { set_error("peer_vec_publish: rank %d placed %d vectors, rank %d placed %d", q, all[q].nvec, rk, P.nvec); return MGCG_ERR_COMM; }
    for (int i = 0; i < P.nvec; ++i) {
        if (rk > 0) P.nb_lo[i] = all[rk - 1].slot[i];
        if (rk < W - 1) P.nb_hi[i] = all[rk + 1].slot[i];
    }
    return MGCG_OK;
}

bool peer_mirror_of(mgcg_ctx *c, const double *ext, Mirror *m)
{
    DistState *D = c->dist;
    PeerState &P = D->peer;
    *m = Mirror();
    if (!P.active || !P.fused) return false;
    const int W = D->world, rk = D->rank;
    for (int i = 0; i < P.nvec; ++i) {
        if (P.vbase[i] != ext) continue;
        m->halo = P.vhalo[i]; m->own = P.vown[i];
        m->ctl = (PeerCtl *)P.arena;
        if (rk > 0) { m->lo = (double *)((char *)P.peer[rk - 1] + P.nb_lo[i].hi_off); m->flag_lo = &((PeerCtl *)P.peer[rk - 1])->xflag[rk]; }
        if (rk < W - 1) { m->hi = (double *)((char *)P.peer[rk + 1] + P.nb_hi[i].lo_off); m->flag_hi = &((PeerCtl *)P.peer[rk + 1])->xflag[rk]; }
        return true;
    }
    return false;
}

bool peer_wait_of(mgcg_ctx *c, GhostWait *g)
{
    DistState *D = c->dist;
    PeerState &P = D->peer;
    *g = GhostWait();
    if (!P.active || !P.fused) return false;
    PeerCtl *me = (PeerCtl *)P.arena;
    g->ctl = me;
    if (D->rank > 0) g->flag_lo = &me->xflag[D->rank - 1];
    if (D->rank < D->world - 1) g->flag_hi = &me->xflag[D->rank + 1];
    g->clocks = P.wait_clocks;
    return true;
}

}  // namespace mgcg

using namespace mgcg;

// Collective micro-benchmark of the loop's exchanges (all ranks call it after mgcg_setup_dist): `reps` back-to-back level-0
// halo exchanges, coarse all-gathers and scalar reductions; out_us[0..2] = microseconds per operation on this rank.
extern "C" int mgcg_dist_microbench(mgcg_ctx *c, int reps, double *out_us)
{
    if (!c || !c->dist || c->dist->L.empty() || !out_us || reps < 1) { set_error("mgcg_dist_microbench: bad argument / not set up"); return MGCG_ERR_INVALID; }
    DistState *D = c->dist;
    MGCG_CUDA(cudaSetDevice(c->device));
    MGCG_CUDA(cudaMemsetAsync(c->d_scalars, 0, sizeof(Scalars), c->stream));
    cudaEvent_t e[4];
    for (int i = 0; i < 4; ++i) MGCG_CUDA(cudaEventCreate(&e[i]));
    DistLevel &L0 = D->L[0];
    for (int w = 0; w < 2; ++w) {      // pass 0 warms up
        MGCG_CUDA(cudaEventRecord(e[0], c->stream));
        for (int i = 0; i < reps; ++i) MGCG_CHECK(dist_halo_exchange(c, D->p, L0.own, L0.lo, L0.hi, L0.halo));
        MGCG_CUDA(cudaEventRecord(e[1], c->stream));
        if (D->coarse) for (int i = 0; i < reps; ++i) MGCG_CHECK(D->peer.active ? peer_allgather(c, D->coarse_b, D->coarse_offs) : MGCG_OK);
        MGCG_CUDA(cudaEventRecord(e[2], c->stream));
        if (D->peer.active) for (int i = 0; i < reps; ++i) MGCG_CHECK(peer_reduce(c, -1, 2, false, 1));
        MGCG_CUDA(cudaEventRecord(e[3], c->stream));
        MGCG_CUDA(cudaStreamSynchronize(c->stream));
    }
    for (int i = 0; i < 3; ++i) { float ms = 0.f; cudaEventElapsedTime(&ms, e[i], e[i + 1]); out_us[i] = 1e3 * ms / reps; }
    for (int i = 0; i < 4; ++i) cudaEventDestroy(e[i]);
    return MGCG_OK;
}
