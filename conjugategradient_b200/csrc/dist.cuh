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
#include <nccl.h>

namespace mgcg {

struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
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
    int first_replicated = 0;               // level index of the replicated tail (0: the whole preconditioner is replicated)
    DevCsr coarse_A;
    double *coarse_b = nullptr;             // all-gathered right-hand side (global length of the first replicated level)
    int coarse_n = 0;
    int coarse_counts[kMaxRanks], coarse_offs[kMaxRanks + 1];
    // CG vectors of level 0
    double *p = nullptr;                    // [lo|own|hi]
    double *r = nullptr;                    // [lo|own|hi] (ghosts used by the restriction)
    double *Ap = nullptr;                   // own
    // scalar reductions: per-rank partials -> all-gather -> fixed rank-order sum (cs:463 resultsDot.Sum())
    double *partial_send = nullptr;         // [2]
    double *partial_all = nullptr;          // [2 * world]
    int halo_exchanges = 0, allgathers = 0; // per-solve accounting
};

// dist.cu
int dist_halo_exchange(mgcg_ctx *c, double *v_ext, int own, int lo, int hi, int halo);
void dist_free(mgcg_ctx *c);

}  // namespace mgcg
