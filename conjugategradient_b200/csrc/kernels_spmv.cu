// kernels_spmv.cu — fp64 CSR SpMV family for sm_100a with fused epilogues (p.Ap dot, residual, Jacobi / Chebyshev sweep).
// Replaces cusparseDcsrmv (Mgcg.cu:17, SimpleConjugateGradient.cu:75,96) and, for the p.Ap epilogue, the
// host-synchronous cublasDdot that follows it (Mgcg.cu:245).  HBM-bound: 12 B per non-zero + 20 B per row on the caller's CSR,
// 9 B per non-zero where the set-up restated the columns as one-byte codes (spmv_code.cu).
//
//  * STREAM  (short rows, stencils): a tile of consecutive rows is a CONTIGUOUS slice of val/col; one thread per row adds its
//    segment in STORAGE order — the same order as SparseMatrix.Multiply (SparseMatrix.cs:78-85), products rounded before the add
//    => bit-identical row sums.  k_spmv_stream: the slice is streamed with coalesced loads and staged as products;
//    k_spmv_tma: val/col (or val/code) slices arrive in a ring of shared-memory stages by 1-D TMA bulk copies on mbarriers.
//  * STENCIL (coded columns in few runs of distances: 5-, 7-, 27-point operators): as k_spmv_tma, and the tile's x operands are
//    staged by TMA too (one contiguous run of x per run of distances), requested by a producer warp: no dependent global load left.
//  * ROWS    (8 .. ~300 per row): TMA-staged slices, L lanes per row, shuffle tree.
//  * VECTOR  (longer rows, unaligned arrays): L lanes per row straight from global memory, shuffle tree.
// All run as persistent grids with a static tile->block map, so the fused reduction is deterministic for a given kernel and grid.
#include "common.cuh"
#include "reduce.cuh"
#include "halo.cuh"
#include <stdlib.h>

namespace mgcg {

constexpr int kSpmvThreads = 256;
constexpr int kChunk = 2048;   // products staged per pass: 16 KB of shared memory

// `wrote` is set when the row's result was also stored into a neighbouring rank's ghost rows (a.mir, sharded solve)
template <int EPI>
__device__ __forceinline__ void epilogue(int row, double s, const SpmvArgs &a, double &dsum, bool &wrote)
{
    double yi;
    if (EPI == EPI_PLAIN) yi = s;
    else if (EPI == EPI_AXPBY) yi = (a.beta == 0.0) ? a.alpha * s : a.alpha * s + a.beta * a.y[row];
    else if (EPI == EPI_DOT) { yi = s; dsum += a.w[row] * s; }
    else if (EPI == EPI_RESID) yi = a.b[row] - s;
    else if (EPI == EPI_JACOBI || EPI == EPI_JACOBI_DOT) {
        const double bi = a.b[row];
        yi = (a.xrow ? a.xrow : a.x)[row] + a.omega * a.dinv[row] * (bi - s);
        if (EPI == EPI_JACOBI_DOT) dsum += bi * yi;
    }
    else {   // EPI_CHEBY, EPI_CHEBY_DOT
        const double bi = a.b[row];
        const double dn = (a.alpha == 0.0 ? 0.0 : a.alpha * a.d_in[row]) + a.omega * a.dinv[row] * (bi - s);
        yi = (a.xrow ? a.xrow : a.x)[row] + dn;
        a.d_out[row] = dn;
        if (EPI == EPI_CHEBY_DOT) dsum += bi * yi;
    }
    a.y[row] = yi;
    if (a.mir.ctl) wrote |= mirror_store(a.mir, row, yi);
}

// The epilogue's operands that do not depend on the row sum, loaded BEFORE the wait for the tile's matrix slice so that their
// latency overlaps it (the prolongator's 3-entry rows leave nothing else to hide a y[row] round trip behind).  Same
// expressions as epilogue<EPI>, operand for operand.
struct EpiPre { double e0, e1, e2, e3; };
template <int EPI>
__device__ __forceinline__ void epi_prefetch(int row, const SpmvArgs &a, EpiPre &p)
{
    p.e0 = 0.0; p.e1 = 0.0; p.e2 = 0.0; p.e3 = 0.0;
    if (EPI == EPI_AXPBY) { if (a.beta != 0.0) p.e0 = a.y[row]; }
    else if (EPI == EPI_DOT) p.e0 = a.w[row];
    else if (EPI == EPI_RESID) p.e0 = a.b[row];
    else if (EPI == EPI_JACOBI || EPI == EPI_JACOBI_DOT) { p.e0 = a.b[row]; p.e1 = (a.xrow ? a.xrow : a.x)[row]; p.e2 = a.dinv[row]; }
    else if (EPI == EPI_CHEBY || EPI == EPI_CHEBY_DOT) {
        p.e0 = a.b[row]; p.e1 = (a.xrow ? a.xrow : a.x)[row]; p.e2 = a.dinv[row];
        if (a.alpha != 0.0) p.e3 = a.d_in[row];
    }
}
template <int EPI>
__device__ __forceinline__ void epilogue_pre(int row, double s, const SpmvArgs &a, double &dsum, const EpiPre &p, bool &wrote)
{
    double yi;
    if (EPI == EPI_PLAIN) yi = s;
    else if (EPI == EPI_AXPBY) yi = (a.beta == 0.0) ? a.alpha * s : a.alpha * s + a.beta * p.e0;
    else if (EPI == EPI_DOT) { yi = s; dsum += p.e0 * s; }
    else if (EPI == EPI_RESID) yi = p.e0 - s;
    else if (EPI == EPI_JACOBI || EPI == EPI_JACOBI_DOT) {
        const double bi = p.e0;
        yi = p.e1 + a.omega * p.e2 * (bi - s);
        if (EPI == EPI_JACOBI_DOT) dsum += bi * yi;
    }
    else {   // EPI_CHEBY, EPI_CHEBY_DOT
        const double bi = p.e0;
        const double dn = (a.alpha == 0.0 ? 0.0 : a.alpha * p.e3) + a.omega * p.e2 * (bi - s);
        yi = p.e1 + dn;
        a.d_out[row] = dn;
        if (EPI == EPI_CHEBY_DOT) dsum += bi * yi;
    }
    a.y[row] = yi;
    if (a.mir.ctl) wrote |= mirror_store(a.mir, row, yi);
}

// red-black half sweeps: is `row` updated by this launch?
__device__ __forceinline__ bool row_active(const SpmvArgs &a, int row)
{
    if (a.color < 0) return true;
    const int ix = row % a.cnx, t = row / a.cnx;
    const int iy = t % a.cny, iz = t / a.cny;
    return ((ix + iy + iz) & 1) == a.color;
}
// epilogue of a row a half sweep leaves alone: x stays, the fused dot still takes b_i x_i
template <int EPI>
__device__ __forceinline__ void epilogue_inactive(const SpmvArgs &a, double &dsum, const EpiPre &p)
{
    if (EPI == EPI_JACOBI_DOT) dsum += p.e0 * p.e1;
}

template <int EPI> struct EpiTraits { static constexpr bool reduces = (EPI == EPI_DOT || EPI == EPI_JACOBI_DOT || EPI == EPI_CHEBY_DOT); };

// ---------------------------------------------------------------------------------------------------------
template <int EPI>
__global__ void __launch_bounds__(kSpmvThreads) k_spmv_stream(SpmvArgs a, int rows_per_tile, int num_tiles,
                                                              ReduceWs ws, Scalars *sc)
{
    pdl_wait(); pdl_launch_dependents();
    if (a.gated && sc->done) return;
    __shared__ double prod[kChunk];
    __shared__ int rp[kSpmvThreads + 1];
    const int t = threadIdx.x;
    const double *__restrict__ val = a.A.val;
    const int *__restrict__ col = a.A.col;
    const double *__restrict__ x = a.x;
    double dsum = 0.0;
    bool wrote = false;
    halo_wait(a.gw, sc);

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
        if (t < nr) {
            if ((EPI == EPI_JACOBI || EPI == EPI_JACOBI_DOT) && !row_active(a, r0 + t)) {
                EpiPre pre;
                epi_prefetch<EPI>(r0 + t, a, pre);
                epilogue_inactive<EPI>(a, dsum, pre);
            } else {
                epilogue<EPI>(r0 + t, acc, a, dsum, wrote);
            }
        }
    }
    if (EpiTraits<EPI>::reduces) {
        double m = 0.0;
        block_reduce<false>(dsum, m);
        grid_reduce_finish<false>(dsum, m, ws, a.fin, sc, a.out);
    }
    halo_signal(a.mir, wrote);
}

// ---------------------------------------------------------------------------------------------------------
// STREAM-TMA: the same tiling, but the contiguous val/col slice of a tile is brought into shared memory by the TMA engine
// (cp.async.bulk, 1-D, completion on an mbarrier) into a ring of stages, one or more tiles ahead of the threads that
// consume it.  The ncu capture of the plain stream kernel shows it latency-bound on the dependent chain
// rowptr -> (val,col) -> x[col] between block barriers (stall long_scoreboard, DRAM 50 %); here the chain is cut: row
// pointers are prefetched a tile ahead in registers, val/col arrive asynchronously, and there is no block barrier in the
// steady state — a thread that has finished its row only arrives on the stage's `empty` mbarrier and moves on.
// One thread per row walks its segment from shared memory in storage order with explicit mul/add roundings.
constexpr int kTmaRows = 256;                 // rows per tile == threads per block
constexpr int kTmaCap = 1920;                 // non-zeros a stage can hold: 2 stages x (8+4) B x 1920 = 45 KB static smem
constexpr int kTmaStagesLong = 2;
// very short rows (prolongators: ~3 per row): a tile is only a few KB, so the same shared memory is cut into twice as
// many stages of half the size — more TMA copies in flight per SM (the ncu capture of the 2-stage kernel on P showed 53 %
// of DRAM peak against 73-75 % on the 7-per-row operator)
constexpr int kTmaCapShort = 960;
constexpr int kTmaStagesShort = 4;
constexpr double kTmaWideMaxAvg = 32.0;       // WIDE variant: 256 rows x 32 x 1.06 x 12 B x 2 stages = 208 KB of dynamic shared memory

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
    unsigned ok;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void tma_load_1d(void *dst, const void *src, unsigned bytes, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct TmaMeta { int k0, k1, direct, pad; };

// kTmaCap == 0: WIDE variant — the stage capacity is a launch parameter and the ring lives in dynamic shared memory (rows of
// 8..32 entries: a 256-row tile is up to ~100 KB, one block per SM).  One thread per row keeps consecutive rows in consecutive
// threads, so on stencil-like matrices (the 27-point operator of C5, the level-1 Galerkin operators) the k-th gathers of a warp fall
// into one 256-byte run of x (the lanes-per-row ROWS kernel spreads a warp's gathers over L runs of 256/L bytes).  Measured, it is
// still behind ROWS there (0.60 against 0.74 of the HBM peak at 256^3): opt-in, see spmv_make_plan.
// kCoded: the matrix carries coded columns (spmv_code.cu) — a stage holds one BYTE per entry instead of an int32 column, and
// col = row + dict[code] is rebuilt from a 256-entry table in shared memory: 9 instead of 12 bytes per entry cross HBM, the entries,
// their order and the x operands are the same, so are the bits of every row sum.  The coded instantiations use the dynamic ring
// (kTmaCap == 0) so that it can hold more stages than the 48 KB of static shared memory allow.
template <int EPI, int kTmaCap, int kTmaStages, bool kCoded>
__global__ void __launch_bounds__(kTmaRows) k_spmv_tma(SpmvArgs a, int num_tiles, int dyn_cap, ReduceWs ws, Scalars *sc)
{
    constexpr bool kDyn = (kTmaCap == 0);
    static_assert(kDyn || !kCoded, "coded columns use the dynamic ring");
    constexpr int kIdxBytes = kCoded ? 1 : 4;            // bytes of a staged column entry
    constexpr int kIdxPad = 16 / kIdxBytes;              // slack entries: a slice starts at a 16-byte boundary of its array
    __shared__ __align__(16) double s_val_st[kDyn ? 1 : kTmaStages][kDyn ? 2 : kTmaCap + 2];
    __shared__ __align__(16) unsigned char s_idx_st[kDyn ? 16 : kTmaStages * (kTmaCap + kIdxPad) * kIdxBytes];
    __shared__ int s_dict[kCoded ? 256 : 1];
    extern __shared__ __align__(16) unsigned char tma_dyn[];
    const int cap = kDyn ? dyn_cap : kTmaCap;                 // multiple of 4 (16 when coded): every stage starts 16-byte aligned
    double *const s_val_b = kDyn ? reinterpret_cast<double *>(tma_dyn) : &s_val_st[0][0];                                   // [stages][cap + 2]
    unsigned char *const s_idx_b = kDyn ? tma_dyn + (size_t)kTmaStages * (cap + 2) * 8 : &s_idx_st[0];                      // [stages][cap + pad]
    const size_t idx_stride = (size_t)(cap + kIdxPad) * kIdxBytes;
    __shared__ __align__(8) unsigned long long s_full[kTmaStages], s_empty[kTmaStages];
    __shared__ TmaMeta s_meta[kTmaStages];
    const int t = threadIdx.x;
    const double *__restrict__ val = a.A.val;
    const int *__restrict__ col = a.A.col;
    const int *__restrict__ rowptr = a.A.rowptr;
    const double *__restrict__ x = a.x;
    const unsigned char *__restrict__ code = a.A.code;
    const int rows = a.A.rows;
    const long long nnz_end = (long long)a.A.nnz;     // entries val/col are known to hold (sub-block views pass the parent's)
    double dsum = 0.0;

    if (t == 0) {
        for (int s = 0; s < kTmaStages; ++s) { mbar_init(&s_full[s], 1); mbar_init(&s_empty[s], kTmaRows); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (kCoded) s_dict[t] = a.A.dict[t];              // kTmaRows == 256 threads
    __syncthreads();

    // the block's tiles: blockIdx.x, blockIdx.x + gridDim.x, ...
    const int my_tiles = (num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

    // producer state (thread 0): issue the slice of local tile `j` into stage j % S
    auto issue = [&](int j) {
        const int s = j % kTmaStages;
        const int tile = blockIdx.x + j * gridDim.x;
        const int r0 = tile * kTmaRows;
        const int nr = min(kTmaRows, rows - r0);
        const int k0 = rowptr[r0], k1 = rowptr[r0 + nr];
        const int ka = k0 & ~1, kc = k0 & ~(kIdxPad - 1);
        const unsigned vb = (unsigned)(((k1 - ka) * 8 + 15) & ~15), cb = (unsigned)(((k1 - kc) * kIdxBytes + 15) & ~15);
        // (the code array is allocated in whole 16-byte pieces with one to spare: its slices never run out of it)
        const bool fits = (k1 - ka) <= cap && (k1 - kc) <= cap && k1 > k0 &&
                          ((long long)ka * 8 + vb <= nnz_end * 8) && (kCoded || (long long)kc * 4 + cb <= nnz_end * 4);
        s_meta[s].k0 = k0; s_meta[s].k1 = k1; s_meta[s].direct = fits ? 0 : 1;
        if (fits) {
            mbar_arrive_expect_tx(&s_full[s], vb + cb);
            tma_load_1d(s_val_b + (size_t)s * (cap + 2), val + ka, vb, &s_full[s]);
            if (kCoded) tma_load_1d(s_idx_b + s * idx_stride, code + kc, cb, &s_full[s]);
            else tma_load_1d(s_idx_b + s * idx_stride, col + kc, cb, &s_full[s]);
        } else {
            mbar_arrive(&s_full[s]);
        }
    };
    if (t == 0)
        for (int j = 0; j < kTmaStages && j < my_tiles; ++j) issue(j);

    // row pointers of this thread's row in the first tile
    int rs_n = 0, re_n = 0;
    {
        const int r = blockIdx.x * kTmaRows + t;
        if (my_tiles > 0 && r < rows) { rs_n = rowptr[r]; re_n = rowptr[r + 1]; }
    }
    // Everything above read only the matrix, which nothing inside the loop writes: under programmatic dependent launch it overlaps the
    // previous kernel's tail.  From here on the vectors and the device scalars of the predecessors are read.
    pdl_wait(); pdl_launch_dependents();
    if (a.gated && sc->done) {       // loop already over: let the requested slices land, then leave
        if (t == 0) for (int j = 0; j < kTmaStages && j < my_tiles; ++j) mbar_wait(&s_full[j], 0u);
        return;
    }
    bool wrote = false;
    halo_wait(a.gw, sc);      // the first tiles' matrix slices are already on their way
    for (int j = 0; j < my_tiles; ++j) {
        const int s = j % kTmaStages;
        const unsigned parity = (unsigned)((j / kTmaStages) & 1);
        const int tile = blockIdx.x + j * gridDim.x;
        const int row = tile * kTmaRows + t;
        const int rs = rs_n, re = re_n;
        // prefetch the next tile's row pointers while this one is processed
        rs_n = re_n = 0;
        if (j + 1 < my_tiles) {
            const int rn = (tile + gridDim.x) * kTmaRows + t;
            if (rn < rows) { rs_n = rowptr[rn]; re_n = rowptr[rn + 1]; }
        }
        EpiPre pre;
        if (row < rows) epi_prefetch<EPI>(row, a, pre);
        const bool active = (EPI == EPI_JACOBI || EPI == EPI_JACOBI_DOT) ? (row < rows && row_active(a, row)) : (row < rows);
        mbar_wait(&s_full[s], parity);
        double acc = 0.0;
        if (active) {
            if (!s_meta[s].direct) {
                const int k0 = s_meta[s].k0;
                const double *sv = s_val_b + (size_t)s * (cap + 2) - (k0 & ~1);
                const int *scl = reinterpret_cast<const int *>(s_idx_b + s * idx_stride) - (k0 & ~3);
                if (kCoded) {
                    const unsigned char *scd = s_idx_b + s * idx_stride - (k0 & ~15);
                    for (int k = rs; k < re; ++k) acc = __dadd_rn(acc, __dmul_rn(sv[k], x[row + s_dict[scd[k]]]));
                } else if (kDyn) {
                    // medium rows: the gathers of four entries are issued together; the sum keeps the storage order
                    int k = rs;
                    for (; k + 4 <= re; k += 4) {
                        const int c0 = scl[k], c1 = scl[k + 1], c2 = scl[k + 2], c3 = scl[k + 3];
                        const double x0 = x[c0], x1 = x[c1], x2 = x[c2], x3 = x[c3];
                        acc = __dadd_rn(acc, __dmul_rn(sv[k], x0));
                        acc = __dadd_rn(acc, __dmul_rn(sv[k + 1], x1));
                        acc = __dadd_rn(acc, __dmul_rn(sv[k + 2], x2));
                        acc = __dadd_rn(acc, __dmul_rn(sv[k + 3], x3));
                    }
                    for (; k < re; ++k) acc = __dadd_rn(acc, __dmul_rn(sv[k], x[scl[k]]));
                } else {
                    for (int k = rs; k < re; ++k) acc = __dadd_rn(acc, __dmul_rn(sv[k], x[scl[k]]));
                }
            } else if (kCoded) {
                for (int k = rs; k < re; ++k) acc = __dadd_rn(acc, __dmul_rn(__ldcs(val + k), x[row + s_dict[__ldcs(code + k)]]));
            } else {
                for (int k = rs; k < re; ++k) acc = __dadd_rn(acc, __dmul_rn(__ldcs(val + k), x[__ldcs(col + k)]));
            }
        }
        mbar_arrive(&s_empty[s]);                 // this thread is done with the stage
        if (t == 0 && j + kTmaStages < my_tiles) {
            mbar_wait(&s_empty[s], parity);       // ... and once every thread is, refill it
            issue(j + kTmaStages);
        }
        if (active) epilogue_pre<EPI>(row, acc, a, dsum, pre, wrote);
        else if (row < rows) epilogue_inactive<EPI>(a, dsum, pre);
    }
    if (EpiTraits<EPI>::reduces) {
        double m = 0.0;
        block_reduce<false>(dsum, m);
        grid_reduce_finish<false>(dsum, m, ws, a.fin, sc, a.out);
    }
    halo_signal(a.mir, wrote);
}

// ---------------------------------------------------------------------------------------------------------
// STENCIL (coded columns whose distances fall into at most kSegMax runs: the 5-, 7- and 27-point operators).  The ncu capture of the
// coded stream kernel showed it latency-bound, not bandwidth-bound: 40 % of the stall samples on the x gathers, 22 % on the wait for
// the next stage — the ring cannot run far enough ahead while its stages are held by threads waiting for gathers.  Here a tile's x
// operands are staged too: the rows r0 .. r0+R-1 of a tile read, for every distance d of the dictionary, the CONTIGUOUS run
// x[r0+d .. r0+d+R-1]; distances within kRunSpan of each other share one run (the 27-point operator: 9 runs of 3 distances), so the
// producer requests one bulk copy per run next to the tile's values and codes.  No thread issues a dependent global load any more:
// rowptr -> (value, code) -> x all arrive through the TMA ring, a row is summed from shared memory in storage order with explicit
// mul/add roundings (the same bits as every other STREAM kernel), and the ring's stages are in flight for almost their whole life.
// A run starts at an index of the same parity as x's own 8-byte misalignment, so that its source is 16-byte aligned.
// Runs that would leave the valid index range of x (tiles at the ends of the matrix, whose rows have no entry at that distance, or
// only some) are not staged: on such a tile the entries of that run fall back to a gather.
// The producer is a warp of its own: with up to fourteen bulk copies per tile behind two row-pointer loads, one consumer thread doing
// the requests between its own rows made the request path the block's critical path (2.1 us per tile whatever the ring depth).  Its
// lanes share the work — lane 0 the values and the barrier's byte count, lane 1 the codes, lane 2 + g run g — and the next tile's row
// pointers are loaded before the wait for the stage to drain.
// R = rows per tile = consumer threads: 256 for rows of up to ~7 entries, 64 up to ~29 (a stage holds kTmaCap entries either way).
template <int EPI, int kStages, int R>
__global__ void __launch_bounds__(R + 32) k_spmv_stencil(SpmvArgs a, int num_tiles, ReduceWs ws, Scalars *sc)
{
    constexpr int cap = kTmaCap;
    constexpr int kRunLen = R + kRunPad;
    extern __shared__ __align__(16) unsigned char st_dyn[];
    const int nseg = a.A.nseg;
    const size_t stage_bytes = (size_t)(cap + 2) * 8 + (cap + 16) + (size_t)nseg * kRunLen * 8;     // values | codes | x runs
    __shared__ __align__(8) unsigned long long s_full[kStages], s_empty[kStages];
    __shared__ TmaMeta s_meta[kStages];       // .pad = bit g set: run g is staged
    __shared__ int s_idx[256];                // code -> position of its operand in the stage's runs (for the tile's row 0)
    __shared__ int s_min[kSegMax];
    const int t = threadIdx.x;
    const double *__restrict__ val = a.A.val;
    const unsigned char *__restrict__ code = a.A.code;
    const int *__restrict__ rowptr = a.A.rowptr;
    const double *__restrict__ x = a.x;
    const int rows = a.A.rows;
    const long long nnz_end = (long long)a.A.nnz;
    const int x_lo = a.x_lo, x_hi = a.x_hi >= 0 ? a.x_hi : a.A.cols;      // valid index range of x
    double dsum = 0.0;

    pdl_wait(); pdl_launch_dependents();
    if (a.gated && sc->done) return;
    if (t == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(&s_full[s], 1); mbar_init(&s_empty[s], R); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // x is 8-byte aligned; a run must start 16-byte aligned: its first index has the parity xpar of x's own misalignment (a shard's
    // shifted view with an odd ghost width), and the position of an operand inside its run moves with it
    const int xpar = (int)((reinterpret_cast<uintptr_t>(x) >> 3) & 1u);
    for (int i = t; i < 256; i += R + 32) {
        int idx = 0;
        if (i < a.A.ndict) {
            const int o = a.A.seg_off[i], g = o >> 10, first = a.A.seg_min[g];
            idx = g * kRunLen + (o & 1023) - (first & 1) + ((first - xpar) & 1);
        }
        s_idx[i] = idx;
    }
    if (t < kSegMax) s_min[t] = t < nseg ? a.A.seg_min[t] : 0;
    bool wrote = false;
    halo_wait(a.gw, sc);             // the x runs are requested below: ghost rows must have landed ...
    if (a.gw.ctl && (t & 31) == 0) asm volatile("fence.proxy.async;" ::: "memory");     // ... and be visible to the bulk copies
    const int my_tiles = (num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    const int full_mask = (1 << nseg) - 1;
    __syncthreads();

    if (t >= R) {
        // ---- producer warp
        const int lane = t - R;
        int k0 = 0, k1 = 0;
        if (my_tiles > 0) {
            const int r0 = blockIdx.x * R;
            k0 = rowptr[r0]; k1 = rowptr[r0 + min(R, rows - r0)];
        }
        for (int j = 0; j < my_tiles; ++j) {
            const int s = j % kStages;
            const int tile = blockIdx.x + j * gridDim.x;
            const int r0 = tile * R;
            int k0n = 0, k1n = 0;
            if (j + 1 < my_tiles) {
                const int rn = (tile + gridDim.x) * R;
                k0n = rowptr[rn]; k1n = rowptr[rn + min(R, rows - rn)];
            }
            if (j >= kStages) mbar_wait(&s_empty[s], (unsigned)(((j / kStages) - 1) & 1));     // the stage's previous tile is consumed
            const int ka = k0 & ~1, kc = k0 & ~15;
            const unsigned vb = (unsigned)(((k1 - ka) * 8 + 15) & ~15), cb = (unsigned)(((k1 - kc) + 15) & ~15);
            // (the code array is allocated in whole 16-byte pieces with one to spare: its slices never run out of it)
            const bool fits = (k1 - ka) <= cap && (k1 - kc) <= cap && k1 > k0 && ((long long)ka * 8 + vb <= nnz_end * 8);
            unsigned char *base = st_dyn + (size_t)s * stage_bytes;
            const int g = lane - 2;
            int s0 = 0;
            bool ok = false;
            if (fits && g >= 0 && g < nseg) {
                s0 = ((r0 + s_min[g] - xpar) & ~1) + xpar;
                ok = s0 >= x_lo && s0 + kRunLen <= x_hi;
            }
            const unsigned mask = __ballot_sync(0xffffffffu, ok) >> 2;
            if (lane == 0) {
                s_meta[s].k0 = k0; s_meta[s].k1 = k1; s_meta[s].direct = fits ? 0 : 1; s_meta[s].pad = (int)mask;
                if (fits) {
                    mbar_arrive_expect_tx(&s_full[s], vb + cb + (unsigned)__popc(mask) * (unsigned)(kRunLen * 8));
                    tma_load_1d(base, val + ka, vb, &s_full[s]);
                } else {
                    mbar_arrive(&s_full[s]);
                }
            } else if (lane == 1) {
                if (fits) tma_load_1d(base + (size_t)(cap + 2) * 8, code + kc, cb, &s_full[s]);
            } else if (ok) {
                double *sx = reinterpret_cast<double *>(base + (size_t)(cap + 2) * 8 + (cap + 16));
                tma_load_1d(sx + g * kRunLen, x + s0, kRunLen * 8, &s_full[s]);
            }
            k0 = k0n; k1 = k1n;
        }
    } else {
        // ---- consumers: one thread per row of the tile
        int rs_n = 0, re_n = 0;
        {
            const int r = blockIdx.x * R + t;
            if (my_tiles > 0 && r < rows) { rs_n = rowptr[r]; re_n = rowptr[r + 1]; }
        }
        for (int j = 0; j < my_tiles; ++j) {
            const int s = j % kStages;
            const unsigned parity = (unsigned)((j / kStages) & 1);
            const int tile = blockIdx.x + j * gridDim.x;
            const int row = tile * R + t;
            const int rs = rs_n, re = re_n;
            rs_n = re_n = 0;
            if (j + 1 < my_tiles) {
                const int rn = (tile + gridDim.x) * R + t;
                if (rn < rows) { rs_n = rowptr[rn]; re_n = rowptr[rn + 1]; }
            }
            EpiPre pre;
            if (row < rows) epi_prefetch<EPI>(row, a, pre);
            const bool active = row < rows;
            mbar_wait(&s_full[s], parity);
            double acc = 0.0;
            if (active) {
                if (!s_meta[s].direct) {
                    const int k0 = s_meta[s].k0, mask = s_meta[s].pad;
                    const unsigned char *base = st_dyn + (size_t)s * stage_bytes;
                    const double *sv = reinterpret_cast<const double *>(base) - (k0 & ~1);
                    const unsigned char *scd = base + (size_t)(cap + 2) * 8 - (k0 & ~15);
                    const double *sx = reinterpret_cast<const double *>(base + (size_t)(cap + 2) * 8 + (cap + 16)) + t;
                    if (mask == full_mask) {
#pragma unroll 4
                        for (int k = rs; k < re; ++k) acc = __dadd_rn(acc, __dmul_rn(sv[k], sx[s_idx[scd[k]]]));
                    } else {
                        for (int k = rs; k < re; ++k) {      // a tile at an end of the matrix: runs that were not staged are gathered
                            const int c = scd[k];
                            const double xv = (mask >> (a.A.seg_off[c] >> 10) & 1) ? sx[s_idx[c]] : x[row + a.A.dict[c]];
                            acc = __dadd_rn(acc, __dmul_rn(sv[k], xv));
                        }
                    }
                } else {
                    for (int k = rs; k < re; ++k) acc = __dadd_rn(acc, __dmul_rn(__ldcs(val + k), x[row + a.A.dict[__ldcs(code + k)]]));
                }
            }
            mbar_arrive(&s_empty[s]);
            if (active) epilogue_pre<EPI>(row, acc, a, dsum, pre, wrote);
        }
    }
    if (EpiTraits<EPI>::reduces) {
        double m = 0.0;
        block_reduce<false>(dsum, m);
        grid_reduce_finish<false>(dsum, m, ws, a.fin, sc, a.out);
    }
    halo_signal(a.mir, wrote);
}

// ---------------------------------------------------------------------------------------------------------
// ROWS-TMA (medium and long rows: 8 .. ~300 entries — the reference's own 160-per-row band, particle matrices, restriction
// operators, coarse Galerkin stencils; cf. the reference's ELL kernel with a local-memory window,
// Mgcg/HandmadeCL/MgcgCL/Mgcg.cl:171-216).  A tile is 256/L consecutive rows, L lanes per row.  The tile's contiguous val/col
// slice is brought into a ring of shared-memory stages by the TMA engine (cp.async.bulk on mbarriers) several tiles ahead, so
// the only global loads the lanes issue themselves are the x gathers (L2-resident for banded matrices); the lanes of a row walk
// its segment from shared memory with stride L and finish with a fixed shuffle tree (deterministic; not the storage order of
// the STREAM kernels).  The plain VECTOR kernel reads val/col through the same dependent chain rowptr -> (val,col) -> x[col]
// and measured 0.69 of the HBM peak on the 79-per-row restriction operator.
// Stage capacity, stage count and resident blocks per SM are launch parameters (spmv_make_plan): the sweep over them on the
// particle matrices (round 2, tools/spmv_sweep.py; the table itself was not kept) showed the kernel short of independent x gathers in flight rather than of
// staged bytes — 2 -> 5 resident blocks per SM took the 35-per-row matrix from 0.42 to 0.81 of the HBM peak with the same ring.
constexpr int kRowsTotalMax = 16384;      // non-zeros the ring may hold: (8 + 4) B x 16384 = 192 KB
constexpr int kRowsMaxStages = 8;
constexpr int kRowsThreads = 256;
// U = entries a lane takes per chunk (its gathers are issued together): 8 when a lane has more than ~6 entries of its row,
// else the plain strided loop unrolled by 4 — the padding of a chunk is predicated off but still costs issue slots (sweep,
// fraction of the HBM peak: 35 per row, L=8: 0.81 plain, 0.78 chunks of 8, 0.71 chunks of 4; 121 per row, L=16: 0.73 plain,
// 0.88 chunks of 8).
template <int EPI, int L, int kRowsUnroll>
__global__ void __launch_bounds__(kRowsThreads) k_spmv_rows(SpmvArgs a, int num_tiles, int cap, int stages, ReduceWs ws, Scalars *sc)
{
    // cap is a multiple of 4: every stage's val slice (cap + 2 doubles) and col slice (cap + 4 ints) starts 16-byte aligned
    extern __shared__ __align__(16) unsigned char rows_smem[];
    double *s_val = reinterpret_cast<double *>(rows_smem);                                   // [stages][cap + 2]
    int *s_col = reinterpret_cast<int *>(rows_smem + (size_t)stages * (cap + 2) * 8);        // [stages][cap + 4]
    unsigned long long *s_full = reinterpret_cast<unsigned long long *>(rows_smem + (size_t)stages * ((cap + 2) * 8 + (cap + 4) * 4));
    unsigned long long *s_empty = s_full + kRowsMaxStages;
    TmaMeta *s_meta = reinterpret_cast<TmaMeta *>(s_empty + kRowsMaxStages);
    constexpr int R = kRowsThreads / L;      // rows per tile
    const int t = threadIdx.x, g = t / L, lane = t % L;
    const double *__restrict__ val = a.A.val;
    const int *__restrict__ col = a.A.col;
    const int *__restrict__ rowptr = a.A.rowptr;
    const double *__restrict__ x = a.x;
    const int rows = a.A.rows;
    const long long nnz_end = (long long)a.A.nnz;
    double dsum = 0.0;

    if (t == 0) {
        for (int s = 0; s < stages; ++s) { mbar_init(&s_full[s], 1); mbar_init(&s_empty[s], kRowsThreads); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int my_tiles = (num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

    auto issue = [&](int j, int s) {
        const int tile = blockIdx.x + j * gridDim.x;
        const int r0 = tile * R;
        const int nr = min(R, rows - r0);
        const int k0 = rowptr[r0], k1 = rowptr[r0 + nr];
        const int ka = k0 & ~1, kc = k0 & ~3;
        const unsigned vb = (unsigned)(((k1 - ka) * 8 + 15) & ~15), cb = (unsigned)(((k1 - kc) * 4 + 15) & ~15);
        const bool fits = (k1 - ka) <= cap && (k1 - kc) <= cap && k1 > k0 &&
                          ((long long)ka * 8 + vb <= nnz_end * 8) && ((long long)kc * 4 + cb <= nnz_end * 4);
        s_meta[s].k0 = k0; s_meta[s].k1 = k1; s_meta[s].direct = fits ? 0 : 1;
        if (fits) {
            mbar_arrive_expect_tx(&s_full[s], vb + cb);
            tma_load_1d(s_val + (size_t)s * (cap + 2), val + ka, vb, &s_full[s]);
            tma_load_1d(s_col + (size_t)s * (cap + 4), col + kc, cb, &s_full[s]);
        } else {
            mbar_arrive(&s_full[s]);
        }
    };
    if (t == 0)
        for (int j = 0; j < stages && j < my_tiles; ++j) issue(j, j);

    int rs_n = 0, re_n = 0;
    {
        const int r = blockIdx.x * R + g;
        if (my_tiles > 0 && r < rows) { rs_n = rowptr[r]; re_n = rowptr[r + 1]; }
    }
    pdl_wait(); pdl_launch_dependents();     // (see k_spmv_tma)
    if (a.gated && sc->done) {
        if (t == 0) for (int j = 0; j < stages && j < my_tiles; ++j) mbar_wait(&s_full[j], 0u);
        return;
    }
    bool wrote = false;
    halo_wait(a.gw, sc);
    int s = 0;
    unsigned parity = 0;
    for (int j = 0; j < my_tiles; ++j) {
        const int tile = blockIdx.x + j * gridDim.x;
        const int row = tile * R + g;
        const int rs = rs_n, re = re_n;
        rs_n = re_n = 0;
        if (j + 1 < my_tiles) {
            const int rn = (tile + gridDim.x) * R + g;
            if (rn < rows) { rs_n = rowptr[rn]; re_n = rowptr[rn + 1]; }
        }
        EpiPre pre;
        if (lane == 0 && row < rows) epi_prefetch<EPI>(row, a, pre);
        mbar_wait(&s_full[s], parity);
        double acc = 0.0;
        if (row < rows) {
            if (!s_meta[s].direct) {
                const int k0 = s_meta[s].k0;
                const double *sv = s_val + (size_t)s * (cap + 2) - (k0 & ~1);
                const int *scl = s_col + (size_t)s * (cap + 4) - (k0 & ~3);
                // chunks of kRowsUnroll entries per lane: all the column indices first, then all the x gathers (independent loads in
                // flight together — the ncu source page of the first version showed the lanes stalled on the gathers' long
                // scoreboard, four at a time), then the products on two accumulators
                if (kRowsUnroll <= 4) {
                    // few entries per lane: the plain strided loop (measured ahead of a padded chunk of 4 on the 35-per-row matrix)
#pragma unroll 4
                    for (int k = rs + lane; k < re; k += L) acc += sv[k] * x[scl[k]];
                } else {
                double acc1 = 0.0;
                for (int kb = rs + lane; kb < re; kb += kRowsUnroll * L) {
                    int cc[kRowsUnroll];
                    double xv[kRowsUnroll];
#pragma unroll
                    for (int u = 0; u < kRowsUnroll; ++u) { const int k = kb + u * L; cc[u] = k < re ? scl[k] : -1; }
#pragma unroll
                    for (int u = 0; u < kRowsUnroll; ++u) xv[u] = cc[u] >= 0 ? x[cc[u]] : 0.0;   // predicated off past the row end:
#pragma unroll                                                                                    // no L1 transaction for the padding
                    for (int u = 0; u < kRowsUnroll; u += 2) {
                        const int ka = kb + u * L, kc = ka + L;
                        acc += (cc[u] >= 0 ? sv[ka] : 0.0) * xv[u];
                        acc1 += (cc[u + 1] >= 0 ? sv[kc] : 0.0) * xv[u + 1];
                    }
                }
                acc += acc1;
                }
            } else {
#pragma unroll 4
                for (int k = rs + lane; k < re; k += L) acc += __ldcs(val + k) * x[__ldcs(col + k)];
            }
        }
        mbar_arrive(&s_empty[s]);
        if (t == 0 && j + stages < my_tiles) {
            mbar_wait(&s_empty[s], parity);
            issue(j + stages, s);
        }
#pragma unroll
        for (int o = L / 2; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o, L);
        if (lane == 0 && row < rows) epilogue_pre<EPI>(row, acc, a, dsum, pre, wrote);
        if (++s == stages) { s = 0; parity ^= 1u; }
    }
    if (EpiTraits<EPI>::reduces) {
        double m = 0.0;
        block_reduce<false>(dsum, m);
        grid_reduce_finish<false>(dsum, m, ws, a.fin, sc, a.out);
    }
    halo_signal(a.mir, wrote);
}

// ---------------------------------------------------------------------------------------------------------
template <int EPI, int L>
__global__ void __launch_bounds__(kSpmvThreads) k_spmv_vector(SpmvArgs a, ReduceWs ws, Scalars *sc)
{
    pdl_wait(); pdl_launch_dependents();
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
    bool wrote = false;
    halo_wait(a.gw, sc);
    for (int row = g; row < rows_padded; row += ngroups) {
        double s = 0.0;
        if (row < rows) {
            const int ks = a.A.rowptr[row], ke = a.A.rowptr[row + 1];
#pragma unroll 4
            for (int k = ks + lane; k < ke; k += L) s += __ldcs(val + k) * x[__ldcs(col + k)];
        }
#pragma unroll
        for (int o = L / 2; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o, L);
        if (lane == 0 && row < rows) epilogue<EPI>(row, s, a, dsum, wrote);
    }
    if (EpiTraits<EPI>::reduces) {
        double m = 0.0;
        block_reduce<false>(dsum, m);
        grid_reduce_finish<false>(dsum, m, ws, a.fin, sc, a.out);
    }
    halo_signal(a.mir, wrote);
}

// ---------------------------------------------------------------------------------------------------------
SpmvPlan spmv_make_plan(const mgcg_ctx *c, int rows, int nnz, int kernel)
{
    SpmvPlan p;
    const double avg = rows > 0 ? (double)nnz / rows : 1.0;
    int cap = c->sm_count * 8;
    if (cap > kMaxPartials) cap = kMaxPartials;
    // AUTO: rows short enough for a 256-row tile to fit a TMA stage -> STREAM (TMA); up to ~300 per row -> ROWS (TMA-staged,
    // L lanes per row); beyond -> VECTOR.  (MGCG_SPMV_NO_ROWS=1 restores the round-1 choice: STREAM up to 48 per row, else VECTOR.)
    // WIDE is opt-in (MGCG_SPMV_WIDE=1): measured BEHIND the ROWS kernel on the 27-point operator (256^3: 3 849 against 4 773 GB/s,
    // tools/spmv_27pt.py) — one resident block of 256 threads per SM keeps too few gathers in flight; it remains for callers that
    // want the storage-order (bit-exact) row sums of the STREAM kernels on medium rows.
    static const bool want_wide = getenv("MGCG_SPMV_WIDE") != nullptr && getenv("MGCG_SPMV_WIDE")[0] == '1';
    const bool wide_ok = want_wide && avg > (double)(kTmaCap - 64) / kTmaRows && avg <= kTmaWideMaxAvg;
    if (kernel == MGCG_SPMV_AUTO) {
        static const bool no_rows = getenv("MGCG_SPMV_NO_ROWS") != nullptr;
        if (avg * kTmaRows <= kTmaCap - 64 || wide_ok) kernel = MGCG_SPMV_STREAM;
        else if (no_rows) kernel = (avg <= 48.0) ? MGCG_SPMV_STREAM : MGCG_SPMV_VECTOR;
        else kernel = (avg <= 300.0) ? MGCG_SPMV_ROWS : MGCG_SPMV_VECTOR;
    }
    p.kernel = kernel;
    if (kernel == MGCG_SPMV_ROWS) {
        // Shape chosen from a round-2 sweep with tools/spmv_sweep.py (table not kept; the chosen shape is what profiles/r02_bench_c3_* measure): what the kernel is short of is
        // independent gathers in flight, not staged bytes — so as many resident blocks as fit (up to 5 per SM), a small ring of 2
        // stages, and L = the smallest power of two that leaves a lane at most ~8 entries of its row.
        // Tuning knobs: MGCG_ROWS_LANES, MGCG_ROWS_STAGES, MGCG_ROWS_CAPF (stage capacity / average tile), MGCG_ROWS_CTAS.
        int L = 4, stages = 2, ctas = 0;
        double capf = 1.3;
        while (L < 32 && avg > 8.0 * L) L <<= 1;
        if (const char *e = getenv("MGCG_ROWS_LANES")) { const int v = atoi(e); if (v == 4 || v == 8 || v == 16 || v == 32) L = v; }
        if (const char *e = getenv("MGCG_ROWS_STAGES")) { const int v = atoi(e); if (v >= 2 && v <= kRowsMaxStages) stages = v; }
        if (const char *e = getenv("MGCG_ROWS_CAPF")) { const double v = atof(e); if (v >= 1.0 && v <= 4.0) capf = v; }
        if (const char *e = getenv("MGCG_ROWS_CTAS")) ctas = atoi(e);
        const double tile = (double)(kRowsThreads / L) * avg;
        long long capn = (long long)(capf * tile) + 16;
        capn = (capn + 3) & ~3LL;
        if (capn < 256) capn = 256;
        while (stages > 2 && capn * stages > kRowsTotalMax) --stages;
        if (capn * stages > kRowsTotalMax) capn = (kRowsTotalMax / stages) & ~3;       // tiles that do not fit take the direct path
        p.lanes = L;
        p.rows_unroll = avg > 6.0 * L ? 8 : 4;
        if (const char *e = getenv("MGCG_ROWS_UNROLL")) { const int v = atoi(e); if (v == 4 || v == 8) p.rows_unroll = v; }
        p.rows_stages = stages;
        p.rows_cap = (int)capn;
        p.rows_smem = (int)((size_t)stages * ((capn + 2) * 8 + (capn + 4) * 4) + (size_t)kRowsMaxStages * 64 + 64);
        const int fit = (int)((224 * 1024) / (p.rows_smem + 1536));
        if (ctas <= 0) ctas = fit < 5 ? fit : 5;
        if (ctas > fit) ctas = fit;
        if (ctas < 1) ctas = 1;
        p.rows_carveout = (int)(((long long)ctas * (p.rows_smem + 1024) * 100 + 228 * 1024 - 1) / (228 * 1024));
        if (p.rows_carveout > 100) p.rows_carveout = 100;
        const int tiles = div_up(rows > 0 ? rows : 1, kRowsThreads / L);
        int gcap = c->sm_count * ctas;
        p.grid = tiles < gcap ? tiles : gcap;
        if (p.grid < 1) p.grid = 1;
        return p;
    }
    // short rows (a 256-row tile fits a TMA stage): the TMA-staged variant of the stream kernel
    p.tma = (kernel == MGCG_SPMV_STREAM && (avg * kTmaRows <= kTmaCap - 64 || wide_ok));
    p.tma_short = p.tma && avg * kTmaRows <= kTmaCapShort - 64;
    if (p.tma && avg * kTmaRows > kTmaCap - 64) {     // WIDE: stage capacity from the average tile, 2 stages in dynamic shared memory
        long long capn = (long long)(avg * kTmaRows * 1.06) + 96;
        capn = (capn + 3) & ~3LL;
        p.tma_wide_cap = (int)capn;
        p.tma_wide_smem = (int)((size_t)kTmaStagesLong * ((capn + 2) * 8 + (capn + 4) * 4));
    }
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

// Dynamic shared memory of the rows kernel: the opt-in limit is raised only as far as a launch needs (per instantiation and
// device), and the L1 / shared-memory split is asked for explicitly — blocks_per_sm x (ring + 1 KB) of shared memory, the rest
// stays L1 for the x gathers.  (A limit set once to the largest ring made every launch run with the smallest L1.)
template <int EPI, int L, int U>
static int rows_prepare_one(mgcg_ctx *c, const SpmvPlan &plan)
{
    static int set_bytes[64] = {0};
    static int set_carve[64] = {0};
    const int d = c->device & 63;
    if (plan.rows_smem > set_bytes[d]) {
        MGCG_CUDA(cudaFuncSetAttribute(k_spmv_rows<EPI, L, U>, cudaFuncAttributeMaxDynamicSharedMemorySize, plan.rows_smem));
        set_bytes[d] = plan.rows_smem;
    }
    int carve = plan.rows_carveout;
    if (const char *e = getenv("MGCG_ROWS_CARVEOUT")) carve = atoi(e);
    if (carve != set_carve[d] && carve > 0) {
        MGCG_CUDA(cudaFuncSetAttribute(k_spmv_rows<EPI, L, U>, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
        set_carve[d] = carve;
    }
    return MGCG_OK;
}
template <int EPI>
static int rows_prepare(mgcg_ctx *c, const SpmvPlan &plan)
{
    const bool u8 = plan.rows_unroll >= 8;
    switch (plan.lanes) {
    case 4: return u8 ? rows_prepare_one<EPI, 4, 8>(c, plan) : rows_prepare_one<EPI, 4, 4>(c, plan);
    case 8: return u8 ? rows_prepare_one<EPI, 8, 8>(c, plan) : rows_prepare_one<EPI, 8, 4>(c, plan);
    case 16: return u8 ? rows_prepare_one<EPI, 16, 8>(c, plan) : rows_prepare_one<EPI, 16, 4>(c, plan);
    default: return u8 ? rows_prepare_one<EPI, 32, 8>(c, plan) : rows_prepare_one<EPI, 32, 4>(c, plan);
    }
}

// Coded columns: the ring lives in dynamic shared memory so that it can be deeper than the int32 kernel's.  A stage is 9 instead of
// 12 bytes per entry; with the int32 kernel's two stages and four resident blocks per SM too few bytes were in flight per SM to
// cover the HBM latency (0.303 ms per fine-level product inside a 256^3 solve against 0.273 ms on int32 columns); three stages x four
// blocks: 0.281 ms.  Stages and resident blocks are tuning knobs (MGCG_CODED_STAGES 2|3, MGCG_CODED_CTAS).
template <int EPI, int STAGES>
static int launch_coded_one(mgcg_ctx *c, const SpmvArgs &a, int tiles, int ctas)
{
    static bool set[64] = {false};
    const int smem = STAGES * ((kTmaCap + 2) * 8 + (kTmaCap + 16));
    const int d = c->device & 63;
    if (!set[d]) {
        MGCG_CUDA(cudaFuncSetAttribute(k_spmv_tma<EPI, 0, STAGES, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        set[d] = true;
    }
    int grid = c->sm_count * ctas;
    if (grid > tiles) grid = tiles;
    launch_k(a.gated, k_spmv_tma<EPI, 0, STAGES, true>, grid, kTmaRows, (size_t)smem, c->stream, a, tiles, kTmaCap, c->ws, c->d_scalars);
    return MGCG_OK;
}
// STENCIL kernel: stages x resident blocks from what fits (tuning knobs MGCG_STENCIL_STAGES 2|3, MGCG_STENCIL_CTAS)
template <int EPI, int STAGES, int R>
static int launch_stencil_one(mgcg_ctx *c, const SpmvArgs &a, int ctas_env)
{
    static int set_bytes[64] = {0};
    const int smem = STAGES * ((kTmaCap + 2) * 8 + (kTmaCap + 16) + a.A.nseg * (R + kRunPad) * 8);
    const int d = c->device & 63;
    if (smem > set_bytes[d]) {
        MGCG_CUDA(cudaFuncSetAttribute(k_spmv_stencil<EPI, STAGES, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        set_bytes[d] = smem;
    }
    int ctas = (224 * 1024) / (smem + 3 * 1024);
    if (ctas > 2048 / (R + 32)) ctas = 2048 / (R + 32);
    if (ctas_env > 0) ctas = ctas_env;
    if (ctas < 1) ctas = 1;
    const int tiles = div_up(a.A.rows, R);
    int grid = c->sm_count * ctas;
    if (grid > tiles) grid = tiles;
    launch_k(a.gated, k_spmv_stencil<EPI, STAGES, R>, grid, R + 32, (size_t)smem, c->stream, a, tiles, c->ws, c->d_scalars);
    return MGCG_OK;
}
template <int EPI>
static int launch_stencil(mgcg_ctx *c, const SpmvArgs &a, int tile_rows)
{
    // (4 stages measured slower than 2 and 3 on both tile shapes and is not instantiated)
    static const int stages = [] { const char *e = getenv("MGCG_STENCIL_STAGES"); const int v = e ? atoi(e) : 0; return v == 3 ? 3 : 2; }();
    static const int ctas_env = [] { const char *e = getenv("MGCG_STENCIL_CTAS"); const int v = e ? atoi(e) : 0; return (v >= 1 && v <= 16) ? v : 0; }();
    if (tile_rows == 64) {
        if (stages == 2) return launch_stencil_one<EPI, 2, 64>(c, a, ctas_env);
        return launch_stencil_one<EPI, 3, 64>(c, a, ctas_env);
    }
    if (stages == 2) return launch_stencil_one<EPI, 2, 256>(c, a, ctas_env);
    return launch_stencil_one<EPI, 3, 256>(c, a, ctas_env);
}

template <int EPI>
static int launch_coded(mgcg_ctx *c, const SpmvArgs &a, int tiles)
{
    static const int stages = [] { const char *e = getenv("MGCG_CODED_STAGES"); const int v = e ? atoi(e) : 0; return v == 2 ? 2 : 3; }();
    static const int ctas_env = [] { const char *e = getenv("MGCG_CODED_CTAS"); const int v = e ? atoi(e) : 0; return (v >= 1 && v <= 8) ? v : 0; }();
    const int fit = (224 * 1024) / (stages * ((kTmaCap + 2) * 8 + (kTmaCap + 16)) + 3 * 1024);
    int ctas = ctas_env > 0 ? ctas_env : fit;
    if (ctas > 6) ctas = 6;
    if (stages == 2) return launch_coded_one<EPI, 2>(c, a, tiles, ctas);
    return launch_coded_one<EPI, 3>(c, a, tiles, ctas);
}

template <int EPI>
static int launch_epi(mgcg_ctx *c, const SpmvPlan &plan, const SpmvArgs &a)
{
    if (a.A.rows <= 0) return MGCG_OK;
    if (a.color >= 0 && (plan.kernel != MGCG_SPMV_STREAM || (EPI != EPI_JACOBI && EPI != EPI_JACOBI_DOT) || a.cnx <= 0 || a.cny <= 0)) {
        set_error("spmv_launch: a red-black half sweep needs a Jacobi epilogue, the STREAM kernels and the grid extents");
        return MGCG_ERR_INVALID;
    }
    const bool aligned = ((((uintptr_t)a.A.val) | ((uintptr_t)a.A.col)) & 15u) == 0;
    // coded columns in few runs: values, codes AND the x runs staged by TMA — unless x is rewritten by this launch
    if (plan.stencil_rows > 0 && a.A.code && a.A.nseg > 0 && aligned && (((uintptr_t)a.x) & 7u) == 0 &&
        (const double *)a.y != a.x && a.color < 0) {
        MGCG_CHECK(launch_stencil<EPI>(c, a, plan.stencil_rows));
        c->launches++;
        MGCG_LAUNCH_CHECK();
        return MGCG_OK;
    }
    if (plan.kernel == MGCG_SPMV_STREAM && plan.tma && aligned) {
        const int tiles = div_up(a.A.rows, kTmaRows);
        if (a.A.code && a.A.dict && spmv_plan_reads_code(plan)) {       // coded, but not eligible for the STENCIL kernel: gathers
            MGCG_CHECK(launch_coded<EPI>(c, a, tiles));
            c->launches++;
            MGCG_LAUNCH_CHECK();
            return MGCG_OK;
        }
        int grid = c->sm_count * 4;
        if (grid > tiles) grid = tiles;
        if (plan.tma_wide_cap > 0) {
            static int set_bytes[64] = {0};
            const int d = c->device & 63;
            if (plan.tma_wide_smem > set_bytes[d]) {
                MGCG_CUDA(cudaFuncSetAttribute(k_spmv_tma<EPI, 0, kTmaStagesLong, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, plan.tma_wide_smem));
                set_bytes[d] = plan.tma_wide_smem;
            }
            int per_sm = (int)((224 * 1024) / (plan.tma_wide_smem + 2048));
            if (per_sm < 1) per_sm = 1;
            if (per_sm > 4) per_sm = 4;
            grid = c->sm_count * per_sm;
            if (grid > tiles) grid = tiles;
            launch_k(a.gated, k_spmv_tma<EPI, 0, kTmaStagesLong, false>, grid, kTmaRows, (size_t)plan.tma_wide_smem, c->stream, a, tiles, plan.tma_wide_cap, c->ws, c->d_scalars);
        }
        else if (plan.tma_short) launch_k(a.gated, k_spmv_tma<EPI, kTmaCapShort, kTmaStagesShort, false>, grid, kTmaRows, 0, c->stream, a, tiles, 0, c->ws, c->d_scalars);
        else launch_k(a.gated, k_spmv_tma<EPI, kTmaCap, kTmaStagesLong, false>, grid, kTmaRows, 0, c->stream, a, tiles, 0, c->ws, c->d_scalars);
    } else if (plan.kernel == MGCG_SPMV_STREAM) {
        int tiles = div_up(a.A.rows, plan.rows_per_tile);
        launch_k(a.gated, k_spmv_stream<EPI>, plan.grid, kSpmvThreads, 0, c->stream, a, plan.rows_per_tile, tiles, c->ws, c->d_scalars);
    } else if (plan.kernel == MGCG_SPMV_ROWS && aligned) {
        MGCG_CHECK(rows_prepare<EPI>(c, plan));
        const int L = plan.lanes;
        const int tiles = div_up(a.A.rows, kRowsThreads / L);
        int grid = plan.grid;
        if (grid > tiles) grid = tiles;
#define MGCG_ROWS_LAUNCH(LL, UU) \
    launch_k(a.gated, k_spmv_rows<EPI, LL, UU>, grid, kRowsThreads, (size_t)plan.rows_smem, c->stream, a, tiles, plan.rows_cap, plan.rows_stages, c->ws, c->d_scalars)
        const bool u8 = plan.rows_unroll >= 8;
        switch (L) {
        case 4: if (u8) MGCG_ROWS_LAUNCH(4, 8); else MGCG_ROWS_LAUNCH(4, 4); break;
        case 8: if (u8) MGCG_ROWS_LAUNCH(8, 8); else MGCG_ROWS_LAUNCH(8, 4); break;
        case 16: if (u8) MGCG_ROWS_LAUNCH(16, 8); else MGCG_ROWS_LAUNCH(16, 4); break;
        default: if (u8) MGCG_ROWS_LAUNCH(32, 8); else MGCG_ROWS_LAUNCH(32, 4); break;
        }
#undef MGCG_ROWS_LAUNCH
    } else {
        switch (plan.lanes) {
        case 4: launch_k(a.gated, k_spmv_vector<EPI, 4>, plan.grid, kSpmvThreads, 0, c->stream, a, c->ws, c->d_scalars); break;
        case 8: launch_k(a.gated, k_spmv_vector<EPI, 8>, plan.grid, kSpmvThreads, 0, c->stream, a, c->ws, c->d_scalars); break;
        case 16: launch_k(a.gated, k_spmv_vector<EPI, 16>, plan.grid, kSpmvThreads, 0, c->stream, a, c->ws, c->d_scalars); break;
        default: launch_k(a.gated, k_spmv_vector<EPI, 32>, plan.grid, kSpmvThreads, 0, c->stream, a, c->ws, c->d_scalars); break;
        }
    }
    c->launches++;
    MGCG_LAUNCH_CHECK();
    return MGCG_OK;
}

bool spmv_plan_reads_code(const SpmvPlan &plan)
{
    return plan.kernel == MGCG_SPMV_STREAM && plan.tma && !plan.tma_short && plan.tma_wide_cap == 0;
}

// tile rows of the STENCIL kernel for a matrix of this shape: a stage holds kTmaCap entries
int spmv_plan_stencil_rows(const SpmvPlan &plan, int rows, int nnz)
{
    const double avg = rows > 0 ? (double)nnz / rows : 0.0;
    if (spmv_plan_reads_code(plan)) return 256;
    if (plan.kernel == MGCG_SPMV_ROWS && avg * 64 <= kTmaCap - 64) return 64;
    return 0;
}

// which kernel family spmv_launch picks for this plan / matrix: 0 stream, 1 TMA-staged stream, 2 lanes-per-row vector, 3 TMA-staged rows,
// 4 TMA-staged stream on coded columns
int spmv_kernel_id(const SpmvPlan &plan, const Csr &A)
{
    const bool aligned = ((((uintptr_t)A.val) | ((uintptr_t)A.col)) & 15u) == 0;
    if (aligned && A.code && A.dict && (plan.stencil_rows > 0 || spmv_plan_reads_code(plan))) return 4;
    if (plan.kernel == MGCG_SPMV_STREAM) return (plan.tma && aligned) ? 1 : 0;
    if (plan.kernel == MGCG_SPMV_ROWS && aligned) return 3;
    return 2;
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
    case EPI_CHEBY: return launch_epi<EPI_CHEBY>(c, plan, a);
    case EPI_CHEBY_DOT: return launch_epi<EPI_CHEBY_DOT>(c, plan, a);
    }
    set_error("spmv_launch: unknown epilogue %d", epi);
    return MGCG_ERR_INVALID;
}

}  // namespace mgcg
