// alloc.cu — size-bucketed block cache under the hierarchy build.
// The build allocates and frees gigabyte-sized transients (sort buffers, A*P) and, on every re-setup, the level operators
// themselves.  Through cudaMallocAsync's pool the same request sequence did NOT reach a steady state: block sizes differ
// from phase to phase, the pool fragments and grows by mapping fresh physical memory at ~75 ms per GB — seen as sporadic
// 90-700 ms stalls inside `R = P^T` / `A*P` of an otherwise 46 ms build (profiles/r01_setup_trace_256.txt).  Here freed
// blocks are kept in per-(device, stream, rounded size) free lists and handed back to the next request of that bucket:
// from the second setup of a given problem on, the build makes no driver allocation at all.
// A bucket miss is served from a SLAB: one large driver allocation per device that blocks are carved from with a bump pointer
// (pool_reserve sizes the next slab from the problem; 512 MiB otherwise).  A driver allocation costs ~0.33 ms whatever its size on
// this box (tools/malloc_cost.py: 100 x 64 MiB = 37 ms) and, once peer access is enabled, is mapped into every peer; a cold
// hierarchy build made ~100-300 of them.  Carved blocks go through the same free lists; a slab returns to the driver when nothing
// carved from it is live and the cache is trimmed.  A cached carved block that is DROPPED (its stream's context went away, or an
// allocation ran out of memory) goes back to its slab as a hole: holes are merged with their neighbours, a hole at the top rewinds
// the bump pointer, and the next carve looks at the holes first — so a slab kept alive by one long-lived block does not leak the
// space of every context that came and went beside it.  Slab sizes: the pool_reserve hint inside a set-up (no 512 MiB floor: a
// 10-row system opens 64 MiB); without a hint 64 MiB doubling to 512 MiB per device, and a request below 1 MiB never opens one.
// Reuse is stream-ordered: a block is only re-issued to work on the stream it was allocated for (the build runs on one
// stream, and a free in host order is a free in stream order).  Sizes are rounded up to 1/8 of their power of two (<= 12.5 %).
#include "common.cuh"
#include <algorithm>
#include <iterator>
#include <map>
#include <vector>
#include <mutex>
#include <tuple>
#include <stdlib.h>

namespace mgcg {

namespace {
struct Block { size_t bytes; int device; cudaStream_t stream; };
typedef std::tuple<int, cudaStream_t, size_t> Key;
std::mutex g_mu;
std::map<void *, Block> g_live;                  // blocks handed out
std::map<Key, std::vector<void *>> g_free;       // blocks waiting for the next request of their bucket
size_t g_cached_bytes = 0;
struct Slab {
    char *base; size_t size, used; int device; long live;
    std::map<size_t, size_t> holes;              // offset -> length of space below `used` that dropped blocks gave back
};
std::vector<Slab> g_slabs;
size_t g_slab_hint = 0;                          // size of the next slab (pool_reserve)
constexpr size_t kSlabDefault = (size_t)512 << 20;   // largest slab opened without a hint
constexpr size_t kSlabFirst = (size_t)64 << 20;      // first slab opened without a hint (doubles per slab of the device)
constexpr size_t kSlabOpenMin = (size_t)1 << 20;     // without a hint, smaller requests do not open a slab

inline size_t carve_size(size_t rb) { return (rb + 255) & ~(size_t)255; }

// [off, off + len) of slab s is free again: merge with adjacent holes; a hole that ends at the bump pointer rewinds it
void slab_give_back(Slab &s, size_t off, size_t len)
{
    auto nx = s.holes.lower_bound(off);
    if (nx != s.holes.begin()) {
        auto pv = std::prev(nx);
        if (pv->first + pv->second == off) { off = pv->first; len += pv->second; s.holes.erase(pv); }
    }
    if (nx != s.holes.end() && off + len == nx->first) { len += nx->second; s.holes.erase(nx); }
    if (off + len == s.used) s.used = off;
    else s.holes[off] = len;
}

// first hole of slab s with room for `need` bytes (taken from its front), or nullptr
void *slab_take_hole(Slab &s, size_t need)
{
    for (auto it = s.holes.begin(); it != s.holes.end(); ++it) {
        if (it->second < need) continue;
        const size_t off = it->first, len = it->second;
        s.holes.erase(it);
        if (len > need) s.holes[off + need] = len - need;
        return s.base + off;
    }
    return nullptr;
}

int slab_of(const void *p)
{
    for (size_t i = 0; i < g_slabs.size(); ++i)
        if ((const char *)p >= g_slabs[i].base && (const char *)p < g_slabs[i].base + g_slabs[i].size) return (int)i;
    return -1;
}

// carve rb bytes (256-byte aligned) for `dev`; nullptr when no slab has room and a new one cannot be made
void *slab_carve(int dev, size_t rb, bool open_new = true)
{
    static const bool off = getenv("MGCG_POOL_NO_SLAB") != nullptr;
    if (off) return nullptr;
    const size_t need = carve_size(rb);
    int on_dev = 0;
    for (Slab &s : g_slabs) {
        if (s.device != dev) continue;
        ++on_dev;
        if (void *p = slab_take_hole(s, need)) { s.live++; return p; }
        if (s.used + need <= s.size) { void *p = s.base + s.used; s.used += need; s.live++; return p; }
    }
    if (!open_new) return nullptr;
    size_t want;
    if (g_slab_hint) want = g_slab_hint;
    else {
        if (need < kSlabOpenMin) return nullptr; // a stray flag or counter outside any set-up: its own (4 KiB) driver allocation
        want = kSlabFirst << (on_dev < 3 ? on_dev : 3);
        if (want > kSlabDefault) want = kSlabDefault;
    }
    if (need > want / 2) return nullptr;         // very large blocks get their own driver allocation
    void *base = nullptr;
    if (cudaMalloc(&base, want) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    Slab s;
    s.base = (char *)base; s.size = want; s.used = need; s.device = dev; s.live = 1;
    g_slabs.push_back(std::move(s));
    return base;
}

// give slabs without live blocks back to the driver (their cached blocks are dropped from the free lists first)
void free_dead_slabs_locked(int device)
{
    for (size_t i = 0; i < g_slabs.size();) {
        Slab &s = g_slabs[i];
        if ((device >= 0 && s.device != device) || s.live > 0) { ++i; continue; }
        for (auto it = g_free.begin(); it != g_free.end();) {
            std::vector<void *> &v = it->second;
            for (size_t k = 0; k < v.size();) {
                if ((char *)v[k] >= s.base && (char *)v[k] < s.base + s.size) { g_cached_bytes -= std::get<2>(it->first); v[k] = v.back(); v.pop_back(); }
                else ++k;
            }
            if (v.empty()) it = g_free.erase(it); else ++it;
        }
        cudaFree(s.base);
        if (i + 1 != g_slabs.size()) g_slabs[i] = std::move(g_slabs.back());
        g_slabs.pop_back();
    }
}

size_t round_up(size_t bytes)
{
    if (bytes <= 4096) return 4096;
    size_t p2 = 4096;
    while ((p2 << 1) <= bytes) p2 <<= 1;
    const size_t step = p2 >> 3;
    return (bytes + step - 1) / step * step;
}

void drop_cached_locked(int device, bool any_stream = true, cudaStream_t stream = nullptr)
{
    std::vector<int> carved_on;                  // devices on which a carved block went back to its slab
    for (auto it = g_free.begin(); it != g_free.end();) {
        if ((device < 0 || std::get<0>(it->first) == device) && (any_stream || std::get<1>(it->first) == stream)) {
            for (void *p : it->second) {
                const int si = slab_of(p);
                if (si < 0) cudaFree(p);
                else {                           // back to the slab as a hole, reusable by any stream of the device
                    slab_give_back(g_slabs[si], (size_t)((char *)p - g_slabs[si].base), carve_size(std::get<2>(it->first)));
                    carved_on.push_back(g_slabs[si].device);
                }
                g_cached_bytes -= std::get<2>(it->first);
            }
            it = g_free.erase(it);
        } else ++it;
    }
    // A cached block is only ever re-issued to its own stream; a hole is not.  Work that stream still has queued on the block must
    // be over before another stream can be handed the space (cudaFree of a driver block waits the same way).  Rare paths only:
    // mgcg_destroy, pool_trim, the out-of-memory retry.
    if (!carved_on.empty()) {
        int cur = 0;
        cudaGetDevice(&cur);
        std::sort(carved_on.begin(), carved_on.end());
        carved_on.erase(std::unique(carved_on.begin(), carved_on.end()), carved_on.end());
        for (int d : carved_on) { cudaSetDevice(d); cudaDeviceSynchronize(); }
        cudaSetDevice(cur);
    }
}
}  // namespace

cudaError_t pool_alloc(void **p, size_t bytes, cudaStream_t st)
{
    *p = nullptr;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const size_t rb = round_up(bytes ? bytes : 1);
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_free.find(Key(dev, st, rb));
    if (it != g_free.end() && !it->second.empty()) {
        *p = it->second.back();
        it->second.pop_back();
        g_cached_bytes -= rb;
        const int si = slab_of(*p);
        if (si >= 0) g_slabs[si].live++;
    } else if ((*p = slab_carve(dev, rb)) == nullptr) {
        e = cudaMalloc(p, rb);
        if (e != cudaSuccess) {          // out of memory: give the cached blocks back and try once more
            cudaGetLastError();
            drop_cached_locked(dev);                               // carved blocks become holes of their slabs ...
            if ((*p = slab_carve(dev, rb, false)) == nullptr) {    // ... which may already have the room
                free_dead_slabs_locked(dev);                       // else whole slabs without a live block go back to the driver
                e = cudaMalloc(p, rb);
                if (e != cudaSuccess) return e;
            }
        }
    }
    g_live[*p] = Block{rb, dev, st};
    return cudaSuccess;
}

// the next slab of this process is at least `bytes` large (mgcg_setup: a multiple of the matrix size)
void pool_reserve(size_t bytes)
{
    std::lock_guard<std::mutex> lk(g_mu);
    g_slab_hint = bytes;
}

// Free a block obtained from pool_alloc (kept for reuse), or anything else (cudaFree).  Safe for work still queued on the
// block's stream: it is only re-issued to that stream.
void pool_release(void *p)
{
    if (!p) return;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        auto it = g_live.find(p);
        if (it != g_live.end()) {
            const Block b = it->second;
            g_live.erase(it);
            const int si = slab_of(p);
            if (si >= 0) g_slabs[si].live--;
            g_free[Key(b.device, b.stream, b.bytes)].push_back(p);
            g_cached_bytes += b.bytes;
            return;
        }
    }
    cudaFree(p);
}

void pool_free(void *p, cudaStream_t) { pool_release(p); }

void pool_trim(int device)
{
    std::lock_guard<std::mutex> lk(g_mu);
    drop_cached_locked(device);
    free_dead_slabs_locked(device);
}

// a context is going away: blocks cached for its stream can never be re-issued
void pool_trim_stream(int device, cudaStream_t st)
{
    std::lock_guard<std::mutex> lk(g_mu);
    drop_cached_locked(device, false, st);
    free_dead_slabs_locked(device);
}

size_t pool_cached_bytes()
{
    std::lock_guard<std::mutex> lk(g_mu);
    return g_cached_bytes;
}

// test hook: {bytes cached in the free lists, slabs of `device`, their total size, bytes below their bump pointers, bytes in holes}
void pool_stats(int device, long long out[5])
{
    std::lock_guard<std::mutex> lk(g_mu);
    out[0] = (long long)g_cached_bytes; out[1] = out[2] = out[3] = out[4] = 0;
    for (const Slab &s : g_slabs) {
        if (device >= 0 && s.device != device) continue;
        out[1]++; out[2] += (long long)s.size; out[3] += (long long)s.used;
        for (const auto &h : s.holes) out[4] += (long long)h.second;
    }
}

}  // namespace mgcg

// test hook (include/mgcg.h): the state of the block cache on `device`; trim != 0 first drops every cached block of the device
extern "C" int mgcg_test_pool_stats(int device, int trim, long long *out5)
{
    if (!out5) { mgcg::set_error("mgcg_test_pool_stats: null argument"); return MGCG_ERR_INVALID; }
    if (trim) mgcg::pool_trim(device);
    mgcg::pool_stats(device, out5);
    return MGCG_OK;
}
