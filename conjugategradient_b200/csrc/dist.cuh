// dist.cuh — one-process-per-GPU sharded MGCG: data structures.
// Replaces the host-threaded ConjugateGradientParallelGpu.Solve (ConjugateGradientParallelGpu.cs:424-565), its
// host-staged halo copies (P2Host/P2Device, Mgcg.cu:88-113) and its host-side partial-dot sums (cs:463,499,525).
//
// Layout.  Rank r owns the contiguous row block [off, off+own) of every DISTRIBUTED level (level 0 = the caller's
// partition; coarser levels = the aggregates of its own rows, so aggregates never cross a shard boundary).  Column
// indices stay GLOBAL on every level (the reference keeps a full-length, globally indexed p for the same reason,
// ConjugateGradientParallelGpu.cs:321): a level vector is stored as [lo ghost | own | hi ghost] and kernels receive the
// pointer shifted by -(off-lo), so x[col] lands in the right slot without rewriting the matrix.  `halo` is the level's
// global half bandwidth (max |col-row|), the generalisation of the reference's minJ/maxJ range rule (Mgcg.cu:83-84,
// cs:397-398): like the reference, only the two neighbouring ranks are exchanged with, and setup fails loudly if a
// column reaches further.  Levels smaller than `replicate_below` rows are REPLICATED: their right-hand side is
// all-gathered and every rank runs the remaining (single-GPU) hierarchy redundantly — the "agglomerated" coarse levels
// of BASELINE.json's north_star, without the scatter back.
#pragma once
#include "mg.cuh"
#include "reduce.cuh"
#include "halo.cuh"
#include <nccl.h>

namespace mgcg {

struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitRankConfig)(ncclComm_t *, int, ncclUniqueId, int, ncclConfig_t *) = nullptr;   // optional (NCCL >= 2.14)
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
int nccl_load(NcclApi **out);   // dlopen("libnccl.so.2") once per process

#define MGCG_NCCL(api, call)                                                                              \
    do {                                                                                                  \
        ncclResult_t r_ = (api)->call;                                                                    \
        if (r_ != ncclSuccess) {                                                                          \
            mgcg::set_error("%s:%d: nccl %s -> %s", __FILE__, __LINE__, #call, (api)->GetErrorString(r_)); \
            return MGCG_ERR_COMM;                                                                         \
        }                                                                                                 \
    } while (0)

constexpr int kMaxRanks = 64;

// ---------------------------------------------------------------------------------------------------------
// peer-memory transport (peer.cu): the loop's exchanges as this library's own kernels over NVLink peer stores
struct PeerPtrs { PeerCtl *ctl[kPeerMax]; };
struct XchgSend { const double *src; double *dst; unsigned long long *flag; int count; int pad; };
struct XchgRecv { const double *stage; double *dst; const unsigned long long *flag; int count; int pad; };
struct XchgDesc { int ns, nr; size_t stage_stride; XchgSend s[kPeerMax]; XchgRecv r[kPeerMax]; };

constexpr int kPeerVecMax = 40;         // sharded vectors that can live in the arena: p, r and three per distributed level
struct PeerVecSlot { size_t lo_off, hi_off; };   // byte offsets (in the owner's arena) of a vector's low- and high-ghost rows

struct PeerState {
    bool fused = false;                 // the level vectors live in the arena: producers store ghost rows directly (halo.cuh)
    int nvec = 0;                       // vectors placed in the arena by the current setup
    size_t vec_off = 0, vec_cap = 0;    // byte offset / capacity of the vector region
    size_t vec_used = 0;
    double *vbase[kPeerVecMax] = {};    // my vectors ([lo|own|hi]) ...
    int vlo[kPeerVecMax] = {}, vown[kPeerVecMax] = {}, vhi[kPeerVecMax] = {}, vhalo[kPeerVecMax] = {};
    PeerVecSlot nb_lo[kPeerVecMax] = {}, nb_hi[kPeerVecMax] = {};   // ... and the ghost slots of the same vector on rank-1 / rank+1
    long long wait_clocks = 8000000000ll;
    bool active = false;                // the loop of the CURRENT setup runs over peer memory
    bool ready = false;                 // arena allocated and every peer mapped (survives a re-setup that fits)
    void *arena = nullptr;              // this rank's arena (cudaMalloc)
    void *peer[kPeerMax] = {};          // every rank's arena as mapped here (peer[rank] == arena)
    bool mapped[kPeerMax] = {};         // opened with cudaIpcOpenMemHandle (to be closed)
    size_t stage_off = 0, stage_cap = 0;   // byte offset of stage[0]; doubles per staging buffer
    int halo_cap = 0;                   // ghost rows per side the staging holds
};

// geometry of one distributed level on every rank (host copy + device copy for the aggregate rule)
struct LevelGeom {
    int world;
    int structured;            // 1: rows are whole z-planes of an nx*ny*nz_r grid per rank
    int nx, ny;                // structured: plane size
    int agg;                   // structured: aggregate edge; unstructured: rows per aggregate (8)
    int cx, cy;                // structured: coarse plane size
    int off[kMaxRanks + 1];    // fine row offsets of the ranks on this level
    int coff[kMaxRanks + 1];   // coarse row offsets (aggregates are numbered rank by rank)
};

struct DistLevel {
    int n_glob = 0, own = 0, off = 0;       // global rows, owned rows, first owned row
    int halo = 0, lo = 0, hi = 0;           // half bandwidth; ghost rows actually present below / above
    Csr A;                                  // own rows, global columns (level 0: the caller's arrays)
    DevCsr Aown;                            // storage for levels >= 1
    DevCsr Pext;                            // prolongator rows of the extended block (setup); P is a view of its own rows
    Csr P;                                  // own rows x global coarse columns
    DevCsr R;                               // owned coarse rows x (lo|own|hi)-local fine columns
    double *dinv = nullptr;                 // own
    double *xa = nullptr, *xb = nullptr;    // [lo|own|hi]
    double *r = nullptr;                    // [lo|own|hi]
    double *b = nullptr;                    // own (levels >= 1; level 0 uses the CG residual, which is [lo|own|hi] too)
    double *xcur = nullptr;
    SpmvPlan planA, planP, planR;
    int cheb = 0;                           // Chebyshev smoother: degree (0: weighted Jacobi)
    double lmax = 0.0;                      // ... Gershgorin bound of D^-1 A over ALL ranks' rows
    double *d = nullptr;                    // ... direction vector (own rows)
    LevelGeom geom;                         // rule that maps this level's rows to the next level's aggregates
    int dims_z = 0;                         // structured: z-planes owned
    bool next_replicated = false;
};

struct DistState {
    NcclApi *api = nullptr;
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    // problem
    int n_glob = 0, off = 0, own = 0;
    std::vector<DistLevel> L;               // distributed levels (at least level 0)
    // replicated tail: a private single-GPU context that owns the all-gathered operator and its hierarchy
    mgcg_ctx *coarse = nullptr;
    Scalars *coarse_scalars_orig = nullptr; // the private context's own scalar block (it runs on the parent's during solves)
    mgcg_ctx *coarse_spare = nullptr;       // the private context of the previous setup, reused by the next one
    int first_replicated = 0;               // level index of the replicated tail (0: the whole preconditioner is replicated)
    DevCsr coarse_A;
    double *coarse_b = nullptr;             // all-gathered right-hand side (global length of the first replicated level)
    int coarse_n = 0;
    int coarse_counts[kMaxRanks], coarse_offs[kMaxRanks + 1];
    // CG vectors of level 0
    double *p = nullptr;                    // [lo|own|hi]
    double *r = nullptr;                    // [lo|own|hi] (ghosts used by the restriction)
    double *Ap = nullptr;                   // own
    bool vec_in_arena = false;              // p, r and the levels' xa / xb / r live in the peer arena (fused halo exchange)
    double *s = nullptr;                    // own; single-reduce CG: s = A p by recurrence (allocated on first use)
    double *cgcg_u = nullptr;               // single-reduce CG: own-rows pointer of the V-cycle's output (fixed across iterations)
    int graph_variant = 0;                  // cg_variant the captured graph was built for
    // scalar reductions: per-rank partials -> all-gather -> fixed rank-order sum (cs:463 resultsDot.Sum())
    double *partial_send = nullptr;         // [2]
    double *partial_all = nullptr;          // [2 * world]
    int *gather_ints = nullptr;             // [2 * world + 2] landing area of the setup's small all-gathers
    int halo_exchanges = 0, allgathers = 0; // per-solve accounting
    PeerState peer;                         // peer-memory transport of the loop (inactive: NCCL)
    // CUDA graph of one sharded iteration (peer transport only)
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t graph_exec = nullptr;
    const double *graph_b = nullptr;
    double *graph_x = nullptr;
    int graph_mode = 0;                     // 1: one iteration per launch; 2: device-side while loop; -2: asked for 2, unsupported
    int graph_halo = 0, graph_gather = 0, graph_nodes = 0;   // per-iteration counts recorded at capture
    // MGCG_DIST_TRACE=1: phase timeline of a solve (eager launches, one CUDA event per phase boundary), printed by rank 0
    bool trace = false;
    std::vector<std::pair<const char *, cudaEvent_t>> trace_ev;
};
inline void dist_trace(mgcg_ctx *c, const char *label)
{
    DistState *D = c->dist;
    if (!D || !D->trace) return;
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) return;
    cudaEventRecord(e, c->stream);
    D->trace_ev.emplace_back(label, e);
}

// solve: per-rank partial sums -> exchange -> every rank adds them in rank order, then the CG bookkeeping
enum DistFin { DF_INIT_PLAIN = 0, DF_INIT_PRE, DF_PAP, DF_RR, DF_RR_RZ, DF_CGCG_INIT, DF_CGCG, DF_CGCG_PLAIN };
__device__ __forceinline__ void dist_finalize(int mode, double s0, double s1, double s2, Scalars *sc)
{
    switch (mode) {
    // single-reduce CG: s0 = w.u, s1 = r.u, s2 = r.r — the iteration's only scalar exchange
    case DF_CGCG_INIT: sc->gamma_new = sc->rz; finalize(FIN_CGCG_INIT, s0, 0.0, sc, nullptr); break;
    case DF_CGCG: sc->rr_pend = s2; sc->rmax = 0.0; sc->gamma_new = s1; finalize(FIN_CGCG_DELTA, s0, 0.0, sc, nullptr); break;
    case DF_CGCG_PLAIN: sc->rr_pend = s2; sc->rmax = 0.0; finalize(FIN_CGCG_DELTA_PLAIN, s0, 0.0, sc, nullptr); break;
    case DF_INIT_PLAIN: finalize(FIN_INIT, s0, 0.0, sc, nullptr); break;
    case DF_INIT_PRE: finalize(FIN_INIT_RR, s0, 0.0, sc, nullptr); finalize(FIN_INIT_RZ, s1, 0.0, sc, nullptr); break;
    case DF_PAP: finalize(FIN_PAP, s0, 0.0, sc, nullptr); break;
    case DF_RR: finalize(FIN_RR, s0, 0.0, sc, nullptr); break;
    case DF_RR_RZ: finalize(FIN_RR_PRECOND, s0, 0.0, sc, nullptr); if (!sc->done) finalize(FIN_RZ, s1, 0.0, sc, nullptr); break;
    }
}

// peer.cu
// collective; leaves peer.active = false if any rank cannot map.  vec_doubles > 0: also reserve that many doubles for the
// sharded vectors (fused halo exchange)
int peer_open(mgcg_ctx *c, int halo_cap, int gather_cap, size_t vec_doubles = 0);
// place a [lo|own|hi] vector in the arena's vector region (zeroed); null when the region is full
double *peer_vec_alloc(mgcg_ctx *c, int lo, int own, int hi, int halo);
// collective: exchange the ghost-slot offsets of the vectors placed so far with the neighbouring ranks
int peer_vec_publish(mgcg_ctx *c);
// producer / consumer descriptors of a vector placed by peer_vec_alloc (ext = its [lo|own|hi] base); false if it is not one
bool peer_mirror_of(mgcg_ctx *c, const double *ext, Mirror *m);
bool peer_wait_of(mgcg_ctx *c, GhostWait *g);
void peer_close(mgcg_ctx *c, bool collective);
int peer_halo_exchange(mgcg_ctx *c, double *v_ext, int own, int lo, int hi, int halo);
int peer_allgather(mgcg_ctx *c, double *full, const int *offs);
int peer_reduce(mgcg_ctx *c, int mode, int k, bool gated, int wait_scale = 1);

// dist.cu
int dist_halo_exchange(mgcg_ctx *c, double *v_ext, int own, int lo, int hi, int halo);
void dist_free(mgcg_ctx *c);

}  // namespace mgcg
