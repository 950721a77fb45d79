// halo.cuh — device side of the peer-memory transport that lives INSIDE the compute kernels (sharded solve, csrc/dist.cu):
//   * a kernel that PRODUCES a sharded vector also stores its boundary rows straight into the neighbouring ranks' ghost rows
//     (plain stores over NVLink into the neighbours' peer-mapped arenas), and its last block publishes the exchange's sequence
//     number in the neighbours' flag words — the halo exchange costs no launch of its own;
//   * the kernel that CONSUMES the ghost rows waits for its neighbours' flags before it touches them.
// Replaces the reference's SyncP = P2Host + P2Device through a host buffer between Solve3 and Solve1
// (Mgcg/cuBlas/Mgcg/ConjugateGradientParallelGpu.cs:384-419, Mgcg/cuBlas/MgcgGpu/Mgcg.cu:88-113).
//
// Write-after-read safety of the direct ghost stores: a rank stores ghost rows of buffer B for exchange s only after it has
// waited for its neighbour's flag of an exchange issued by a kernel that FOLLOWS the neighbour's last reader of B in stream
// order, or after a scalar reduction / all-gather (full barriers) that follows it — the loop's kernel order guarantees one of
// the two for every buffer (csrc/dist.cu, comment at dist_vcycle).
#pragma once
#include "common.cuh"

namespace mgcg {

constexpr int kPeerMax = 16;            // ranks of one NVSwitch domain served by the peer transport (NCCL beyond that)

struct PeerCtl {                        // head of every rank's arena; remote ranks write xflag / rflag / rslot
    unsigned long long xflag[kPeerMax]; // xflag[q]: sequence number of the last exchange whose data from rank q has landed
    unsigned long long rflag[kPeerMax]; // the same for the scalar reductions
    double rslot[2][kPeerMax][4];       // [parity][sender][value] (up to 3 sums per reduction)
    unsigned long long xseq, rseq;      // local: exchanges / reductions completed (device-resident so graphs replay)
    unsigned int ticket1, ticket2, pad[2];
};

__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// true when the flag reached `seq`; false after `limit` clocks
__device__ __forceinline__ bool wait_flag_clocks(const unsigned long long *flag, unsigned long long seq, long long limit)
{
    const long long t0 = clock64();
    while (ld_acquire_sys(flag) < seq) {
        if (clock64() - t0 > limit) return false;
        __nanosleep(32);
    }
    return true;
}

// value of own row `row` just written to the vector: mirror it if it is a boundary row.  Returns true when it stored remotely.
__device__ __forceinline__ bool mirror_store(const Mirror &m, int row, double v)
{
    bool wrote = false;
    if (m.lo && row < m.halo) { m.lo[row] = v; wrote = true; }
    if (m.hi && row >= m.own - m.halo) { m.hi[row - (m.own - m.halo)] = v; wrote = true; }
    return wrote;
}

// End of a producing kernel, called by ALL threads of every block: the block that finishes last publishes the exchange.
__device__ __forceinline__ void halo_signal(const Mirror &m, bool wrote)
{
    if (!m.ctl) return;
    if (wrote) __threadfence_system();          // my remote stores are visible system-wide before my block's ticket
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned int k = atomicAdd(&m.ctl->ticket1, 1u);
        if (k == gridDim.x - 1) {
            m.ctl->ticket1 = 0u;
            __threadfence_system();
            const unsigned long long seq = ld_volatile_u64(&m.ctl->xseq) + 1ull;
            if (m.flag_lo) st_release_sys(m.flag_lo, seq);
            if (m.flag_hi) st_release_sys(m.flag_hi, seq);
            m.ctl->xseq = seq;                  // read by the kernels that follow on this stream
        }
    }
}

// Start of a consuming kernel, called by ALL threads of every block.  The exchange waited for is the latest one this rank has
// itself issued (every rank runs the same sequence of exchanges).
__device__ __forceinline__ void halo_wait(const GhostWait &g, Scalars *sc)
{
    if (!g.ctl) return;
    if (threadIdx.x == 0) {
        const unsigned long long seq = ld_volatile_u64(&g.ctl->xseq);
        bool ok = true;
        if (g.flag_lo && !wait_flag_clocks(g.flag_lo, seq, g.clocks)) ok = false;
        if (g.flag_hi && !wait_flag_clocks(g.flag_hi, seq, g.clocks)) ok = false;
        if (!ok) { sc->status = MGCG_ERR_COMM; sc->done = 1; }
    }
    __syncthreads();
}

}  // namespace mgcg
