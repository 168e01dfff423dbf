// onesweep.cuh — stable LSD radix sort of 64-bit words in one read + one write per 8-bit digit pass.
//
// The three-kernel pass of sortscan.cuh (block histogram -> scan of all block histograms -> scatter) reads the
// keys twice per pass and spends a third of its time in the small scan launches.  Here:
//   * ONE up-front pass computes the global digit histograms of ALL passes (they do not depend on the order of
//     the keys), fused into whatever kernel produces the keys (os_hist_accumulate);
//   * each pass is ONE kernel: a tile of 2048 keys is ranked in shared memory exactly as in sortscan.cuh, and the
//     tile's position inside every digit's run comes from a decoupled look-back over the preceding tiles' counts
//     (a tile publishes its per-digit count, then the inclusive prefix, as one 32-bit word per digit: value | flag),
//     so nothing but the keys is read twice.  Tiles take their index from an atomic counter in the order they
//     start running, hence a tile only ever waits for tiles that are already running or done — no deadlock
//     whatever the scheduler does; the polling loop is bounded anyway and raises a flag instead of hanging.
// Payloads ride in the low bits of the 64-bit word (arrival index, small fields): the sort only looks at
// [begin_bit, end_bit).  n < 2^30 (counts share their word with two flag bits).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "igs_internal.cuh"

namespace {

constexpr int kOsThreads = 256;
constexpr int kOsItems = 8;
constexpr int kOsTile = kOsThreads * kOsItems;      // 2048 keys per tile
constexpr int kOsWarps = kOsThreads / 32;
constexpr int kOsBins = 256;
constexpr int kOsSpan = kOsItems * 32;              // keys per warp
constexpr int kOsMaxPasses = 8;
constexpr uint32_t kOsFlagAgg = 1u << 30, kOsFlagInc = 1u << 31, kOsMask = (1u << 30) - 1u;
constexpr uint32_t kOsMaxPolls = 1u << 24;          // ~seconds; a healthy look-back needs a handful

struct OsPlan {
    int passes;                  // number of 8-bit digit passes
    int shift[kOsMaxPasses];     // bit position of each pass's digit
    int bits[kOsMaxPasses];      // width of each digit (the top one may be narrower)
};

inline OsPlan os_plan(int begin_bit, int end_bit) {
    OsPlan p;
    p.passes = 0;
    for (int b = begin_bit; b < end_bit && p.passes < kOsMaxPasses; b += 8) {
        p.shift[p.passes] = b;
        p.bits[p.passes] = end_bit - b < 8 ? end_bit - b : 8;
        ++p.passes;
    }
    return p;
}

// Scratch of one sort: ghist [passes][256] (u32, zeroed), counters [passes] (zeroed), status [passes][tiles][256]
// (zeroed), one error word (zeroed).  All in one allocation so that one memset prepares a sort.
struct OsScratch {
    uint32_t* base = nullptr;
    uint32_t n_tiles = 0;
    int passes = 0;
    __host__ __device__ uint32_t* ghist(int p) const { return base + (size_t)p * kOsBins; }
    __host__ __device__ uint32_t* counter(int p) const { return base + (size_t)passes * kOsBins + p; }
    __host__ __device__ uint32_t* error() const { return base + (size_t)passes * kOsBins + kOsMaxPasses; }
    __host__ __device__ uint32_t* status(int p) const {
        return base + (size_t)passes * kOsBins + kOsMaxPasses + 8 + (size_t)p * n_tiles * kOsBins;
    }
    static size_t words(uint32_t n, int passes) {
        const size_t tiles = (n + kOsTile - 1) / kOsTile;
        return (size_t)passes * kOsBins + kOsMaxPasses + 8 + (size_t)passes * tiles * kOsBins;
    }
};

// ---- global digit histograms of all passes -----------------------------------------------------
// Per-warp tables in shared memory ([warp][pass][256]); call os_hist_flush at the end of the block.
__device__ __forceinline__ void os_hist_clear(uint32_t* sh, int passes) {
    for (int i = threadIdx.x; i < kOsWarps * passes * kOsBins; i += blockDim.x) sh[i] = 0;
}
__device__ __forceinline__ void os_hist_accumulate(uint32_t* sh, const OsPlan& plan, uint64_t key) {
    uint32_t* mine = sh + (threadIdx.x >> 5) * plan.passes * kOsBins;
#pragma unroll 1
    for (int p = 0; p < plan.passes; ++p)
        atomicAdd(mine + p * kOsBins + ((uint32_t)(key >> plan.shift[p]) & ((1u << plan.bits[p]) - 1u)), 1u);
}
__device__ __forceinline__ void os_hist_flush(const uint32_t* sh, const OsPlan& plan, const OsScratch& sc) {
    for (int i = threadIdx.x; i < plan.passes * kOsBins; i += blockDim.x) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < kOsWarps; ++w) t += sh[w * plan.passes * kOsBins + i];
        if (t) atomicAdd(sc.ghist(i / kOsBins) + (i % kOsBins), t);
    }
}
inline size_t os_hist_smem_bytes(int passes) { return (size_t)kOsWarps * passes * kOsBins * sizeof(uint32_t); }

// stand-alone histogram kernel for keys that already exist
__global__ void __launch_bounds__(kOsThreads)
k_os_hist(const uint64_t* __restrict__ keys, uint32_t n, OsPlan plan, OsScratch sc) {
    extern __shared__ uint32_t os_sh_hist[];
    __shared__ OsPlan sh_plan;
    if (threadIdx.x == 0) sh_plan = plan;
    os_hist_clear(os_sh_hist, plan.passes);
    __syncthreads();
    for (uint32_t i = blockIdx.x * kOsThreads + threadIdx.x; i < n; i += gridDim.x * kOsThreads)
        os_hist_accumulate(os_sh_hist, sh_plan, __ldg(keys + i));
    __syncthreads();
    os_hist_flush(os_sh_hist, sh_plan, sc);
}

// exclusive scan of each pass's 256 global counts, in place (one block, one warp per pass)
__global__ void __launch_bounds__(32 * kOsMaxPasses)
k_os_scan_hist(OsScratch sc) {
    const int p = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (p >= sc.passes) return;
    uint32_t* h = sc.ghist(p);
    uint32_t v[kOsBins / 32], sum = 0;
#pragma unroll
    for (int k = 0; k < kOsBins / 32; ++k) { v[k] = sum; sum += h[lane * (kOsBins / 32) + k]; }
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += y;
    }
    const uint32_t excl = inc - sum;
#pragma unroll
    for (int k = 0; k < kOsBins / 32; ++k) h[lane * (kOsBins / 32) + k] = v[k] + excl;
}

struct OsSmem {
    uint64_t keys[kOsTile];
    uint32_t cnt[kOsWarps][kOsBins];   // per warp and digit: count, then start inside the digit's run of the tile
    uint32_t tile_base[kOsBins];       // start of the digit's run inside the tile
    uint32_t gbase[kOsBins];           // start of this tile's part of the digit's run in the output
    uint32_t tile_id;
};

// One digit pass: stable scatter of keys_in into keys_out.
__global__ void __launch_bounds__(kOsThreads, 6)
k_os_pass(const uint64_t* __restrict__ keys_in, uint64_t* __restrict__ keys_out, uint32_t n, int pass, int shift,
          uint32_t bins, OsScratch sc) {
    extern __shared__ __align__(16) unsigned char os_smem_raw[];
    OsSmem& s = *reinterpret_cast<OsSmem*>(os_smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) s.tile_id = atomicAdd(sc.counter(pass), 1u);
    for (uint32_t d = threadIdx.x; d < kOsWarps * kOsBins; d += kOsThreads) (&s.cnt[0][0])[d] = 0;
    __syncthreads();
    const uint32_t tile = s.tile_id;
    uint32_t* my_cnt = s.cnt[warp];
    const uint32_t tile0 = tile * kOsTile;
    const uint32_t wbase = tile0 + warp * kOsSpan;
    uint64_t key[kOsItems];
    uint32_t rank[kOsItems];
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kOsItems; ++r) {
        const uint32_t i = wbase + r * 32 + lane;
        const bool valid = i < n;
        key[r] = valid ? __ldg(keys_in + i) : 0;
        const uint32_t digit = valid ? ((uint32_t)(key[r] >> shift) & (bins - 1)) : (kOsBins + lane);
        const uint32_t peers = __match_any_sync(0xffffffffu, digit);
        const uint32_t before = valid ? my_cnt[digit] : 0;
        __syncwarp();
        if (valid && (peers & lt_mask) == 0) my_cnt[digit] = before + __popc(peers);
        __syncwarp();
        rank[r] = before + __popc(peers & lt_mask);
    }
    __syncthreads();
    // one thread per digit: order the warps and publish the tile's count at once, so that later tiles can add it
    // up while this one is still busy
    const uint32_t d = threadIdx.x;                // kOsThreads == kOsBins
    volatile uint32_t* st = sc.status(pass);
    uint32_t run = 0;
    {
#pragma unroll
        for (int w = 0; w < kOsWarps; ++w) {
            const uint32_t c = s.cnt[w][d];
            s.cnt[w][d] = run;
            run += c;
        }
        s.tile_base[d] = run;                      // totals; scanned below
        if (d < bins) st[(size_t)tile * kOsBins + d] = run | (tile == 0 ? kOsFlagInc : kOsFlagAgg);
    }
    __syncthreads();
    if (warp == 0) {                               // exclusive scan of the 256 totals by one warp
        uint32_t v[kOsBins / 32], sum = 0;
#pragma unroll
        for (int k = 0; k < kOsBins / 32; ++k) { v[k] = sum; sum += s.tile_base[lane * (kOsBins / 32) + k]; }
        uint32_t inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += y;
        }
        const uint32_t excl = inc - sum;
#pragma unroll
        for (int k = 0; k < kOsBins / 32; ++k) s.tile_base[lane * (kOsBins / 32) + k] = v[k] + excl;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kOsItems; ++r) {           // lay the tile out by digit in shared memory
        const uint32_t i = wbase + r * 32 + lane;
        if (i < n) {
            const uint32_t digit = (uint32_t)(key[r] >> shift) & (bins - 1);
            s.keys[s.tile_base[digit] + my_cnt[digit] + rank[r]] = key[r];
        }
    }
    {   // look back: sum the counts of the preceding tiles until one of them knows its inclusive prefix.  Four
        // predecessors are polled at a time (independent loads), then consumed in order.  (Polling 8 .. 32 at a time
        // shortens the chain of a wave of tiles that starts at once but costs more in traffic and registers than it
        // saves: 95 -> 210 us per 10M-key pass.)
        uint32_t excl = 0;
        if (d < bins && tile != 0) {
            uint32_t polls = 0;
            uint32_t t = tile;                     // next predecessor to consume is t - 1
            bool done = false;
            while (!done && t > 0) {
                uint32_t v[4];
                const uint32_t m = t < 4 ? t : 4;
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k) v[k] = k < m ? st[(size_t)(t - 1 - k) * kOsBins + d] : 0u;
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k) {
                    if (done || k >= m) break;
                    if ((v[k] & (kOsFlagAgg | kOsFlagInc)) == 0) break;         // not published yet: poll again from here
                    excl += v[k] & kOsMask;
                    --t;
                    if (v[k] & kOsFlagInc) done = true;
                }
                if (++polls > kOsMaxPolls) { *sc.error() = 1u; done = true; }
            }
            st[(size_t)tile * kOsBins + d] = ((excl + run) & kOsMask) | kOsFlagInc;
        }
        s.gbase[d] = d < bins ? sc.ghist(pass)[d] + excl : 0;
    }
    __syncthreads();
    const uint32_t tile_n = min((uint32_t)kOsTile, n - tile0);
    for (uint32_t j = threadIdx.x; j < tile_n; j += kOsThreads) {
        const uint64_t k = s.keys[j];
        const uint32_t digit = (uint32_t)(k >> shift) & (bins - 1);
        keys_out[s.gbase[digit] + (j - s.tile_base[digit])] = k;
    }
}
static_assert(kOsThreads == kOsBins, "one thread per digit value in the look-back step");

// Prepares the scratch of a sort (one memset).  The caller then fills the global histograms — k_os_hist, or its
// own key-producing kernel through os_hist_* — and calls os_sort_passes.
inline int os_prepare(igs_engine* e, OsScratch* sc, uint32_t* words, uint32_t n, const OsPlan& plan, cudaStream_t st) {
    sc->base = words;
    sc->n_tiles = (n + kOsTile - 1) / kOsTile;
    sc->passes = plan.passes;
    IGS_CUDA(e, cudaMemsetAsync(words, 0, OsScratch::words(n, plan.passes) * sizeof(uint32_t), st));
    return IGS_OK;
}

// Runs the digit passes (global histograms already accumulated in sc).  keys[0] holds the input; *cur_out = index
// of the buffer with the result.  *d_error (device word inside sc) is non-zero afterwards if a look-back gave up.
inline int os_sort_passes(igs_engine* e, uint64_t* keys[2], uint32_t n, const OsPlan& plan, const OsScratch& sc,
                          cudaStream_t st, int* cur_out) {
    IGS_CUDA(e, cudaFuncSetAttribute(k_os_pass, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(OsSmem)));
    k_os_scan_hist<<<1, 32 * kOsMaxPasses, 0, st>>>(sc);
    IGS_LAUNCHED(e);
    int cur = 0;
    for (int p = 0; p < plan.passes; ++p) {
        k_os_pass<<<sc.n_tiles, kOsThreads, sizeof(OsSmem), st>>>(keys[cur], keys[cur ^ 1], n, p, plan.shift[p],
                                                                  1u << plan.bits[p], sc);
        IGS_LAUNCHED(e);
        cur ^= 1;
    }
    *cur_out = cur;
    return IGS_OK;
}

// Whole sort of existing keys on bits [begin_bit, end_bit).
inline int os_sort_u64(igs_engine* e, uint64_t* keys[2], uint32_t n, int begin_bit, int end_bit, uint32_t* words,
                       cudaStream_t st, int* cur_out, OsScratch* sc_out = nullptr) {
    const OsPlan plan = os_plan(begin_bit, end_bit);
    OsScratch sc;
    int rc = os_prepare(e, &sc, words, n, plan, st);
    if (rc != IGS_OK) return rc;
    IGS_CUDA(e, cudaFuncSetAttribute(k_os_hist, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)os_hist_smem_bytes(plan.passes)));
    const uint32_t blocks = std::min<uint32_t>((n + kOsThreads - 1) / kOsThreads, (uint32_t)e->sm_count * 4);
    k_os_hist<<<blocks, kOsThreads, os_hist_smem_bytes(plan.passes), st>>>(keys[0], n, plan, sc);
    IGS_LAUNCHED(e);
    if (sc_out) *sc_out = sc;
    return os_sort_passes(e, keys, n, plan, sc, st, cur_out);
}

}  // namespace
