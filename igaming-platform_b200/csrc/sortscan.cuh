// sortscan.cuh — building blocks shared by the index build (k_velocity.cu) and the linked-accounts
// stages (k_links.cu): stable LSD radix sort of (u64 key, u32 value) pairs, u32 exclusive scan, and the
// engine's grow-only scratch pool.  Internal header: everything has internal linkage per translation unit.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "igs_internal.cuh"

namespace {

constexpr int kThreads = 256;

// ------------------------------------------------------------------------------------------------
// LSD radix sort pass: histogram -> exclusive scan (bin-major) -> stable scatter
// ------------------------------------------------------------------------------------------------
// Digits of at most 8 bits on tiles of 2048 keys: a tile then holds ~8 keys per digit value, so after the
// tile has been ordered by digit in shared memory, neighbouring threads write neighbouring global addresses
// (runs of ~64 bytes of keys: whole 32-byte sectors) instead of one sector per key.  (10-bit digits scattered
// straight from registers — the first version — sent every warp's 32 keys to ~32 different sectors and ran at
// a quarter of the HBM roofline.  4096-key tiles coalesce even better but fit only 3 blocks per SM: 104 us per
// pass against 82 us; 1024-key tiles: 84 us and a costlier histogram.)
constexpr int kSortItems = 8;
constexpr int kSortTile = kThreads * kSortItems;   // keys per block
constexpr int kSortWarps = kThreads / 32;
constexpr int kSortMaxBits = 8;
constexpr int kSortBins = 1 << kSortMaxBits;
constexpr int kSortSpan = kSortItems * 32;         // keys per warp

__global__ void __launch_bounds__(kThreads)
k_radix_hist(const uint64_t* __restrict__ keys, uint32_t n, int shift, uint32_t bins,
             uint32_t* __restrict__ hist_g, uint32_t nblocks) {
    __shared__ uint32_t sh_hist[kSortWarps][kSortBins];     // per-warp tables: less contention on hot digits
    for (uint32_t d = threadIdx.x; d < kSortWarps * kSortBins; d += kThreads) (&sh_hist[0][0])[d] = 0;
    __syncthreads();
    const int warp = threadIdx.x >> 5;
    const uint32_t base = blockIdx.x * kSortTile;
#pragma unroll
    for (int r = 0; r < kSortItems; ++r) {
        const uint32_t i = base + r * kThreads + threadIdx.x;
        if (i < n) atomicAdd(&sh_hist[warp][(uint32_t)(__ldg(keys + i) >> shift) & (bins - 1)], 1u);
    }
    __syncthreads();
    for (uint32_t d = threadIdx.x; d < bins; d += kThreads) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) t += sh_hist[w][d];
        hist_g[(size_t)d * nblocks + blockIdx.x] = t;
    }
}

struct SortSmem {
    uint64_t keys[kSortTile];
    uint32_t vals[kSortTile];
    uint32_t cnt[kSortWarps][kSortBins];   // per warp and digit: count, then start inside the digit's run
    uint32_t tile_base[kSortBins];         // start of the digit's run inside the tile
    uint32_t gbase[kSortBins];             // start of this tile's part of the digit's run in the output
};

// Stable scatter.  Warp w owns a contiguous span of the tile (kSortSpan keys) and walks it in rounds of
// 32 keys; __match_any_sync ranks equal digits inside a round, a per-warp counter table carries the rank
// across rounds, a per-digit scan over the 8 warps orders the warps.  The tile is then laid out by digit in
// shared memory and written out with consecutive threads on consecutive slots.
__global__ void __launch_bounds__(kThreads)
k_radix_scatter(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, uint32_t n, int shift,
                uint32_t bins, const uint32_t* __restrict__ hist_scan_g, uint32_t nblocks) {
    extern __shared__ __align__(16) unsigned char sort_smem_raw[];
    SortSmem& s = *reinterpret_cast<SortSmem*>(sort_smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (uint32_t d = threadIdx.x; d < kSortWarps * kSortBins; d += kThreads) (&s.cnt[0][0])[d] = 0;
    __syncthreads();
    uint32_t* my_cnt = s.cnt[warp];
    const uint32_t tile0 = blockIdx.x * kSortTile;
    const uint32_t wbase = tile0 + warp * kSortSpan;
    uint64_t key[kSortItems];
    uint32_t rank[kSortItems];
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kSortItems; ++r) {
        const uint32_t i = wbase + r * 32 + lane;
        const bool valid = i < n;
        key[r] = valid ? __ldg(keys_in + i) : 0;
        const uint32_t digit = valid ? ((uint32_t)(key[r] >> shift) & (bins - 1)) : (kSortBins + lane);
        const uint32_t peers = __match_any_sync(0xffffffffu, digit);
        const uint32_t before = valid ? my_cnt[digit] : 0;
        __syncwarp();
        if (valid && (peers & lt_mask) == 0) my_cnt[digit] = before + __popc(peers);
        __syncwarp();
        rank[r] = before + __popc(peers & lt_mask);
    }
    __syncthreads();
    // per digit: exclusive scan over the warps, the tile total, this tile's base in the output
    {
        const uint32_t d = threadIdx.x;            // kThreads == kSortBins
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) {
            const uint32_t c = s.cnt[w][d];
            s.cnt[w][d] = run;
            run += c;
        }
        s.tile_base[d] = run;                      // totals; scanned below
        s.gbase[d] = d < bins ? __ldg(hist_scan_g + (size_t)d * nblocks + blockIdx.x) : 0;
    }
    __syncthreads();
    if (warp == 0) {                               // exclusive scan of the 256 totals by one warp
        uint32_t v[kSortBins / 32], sum = 0;
#pragma unroll
        for (int k = 0; k < kSortBins / 32; ++k) { v[k] = sum; sum += s.tile_base[lane * (kSortBins / 32) + k]; }
        uint32_t inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += y;
        }
        const uint32_t excl = inc - sum;
#pragma unroll
        for (int k = 0; k < kSortBins / 32; ++k) s.tile_base[lane * (kSortBins / 32) + k] = v[k] + excl;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kSortItems; ++r) {
        const uint32_t i = wbase + r * 32 + lane;
        if (i < n) {
            const uint32_t digit = (uint32_t)(key[r] >> shift) & (bins - 1);
            const uint32_t pos = s.tile_base[digit] + my_cnt[digit] + rank[r];
            s.keys[pos] = key[r];
            s.vals[pos] = __ldg(vals_in + i);
        }
    }
    __syncthreads();
    const uint32_t tile_n = min((uint32_t)kSortTile, n - tile0);
    for (uint32_t j = threadIdx.x; j < tile_n; j += kThreads) {
        const uint64_t k = s.keys[j];
        const uint32_t digit = (uint32_t)(k >> shift) & (bins - 1);
        const uint32_t out = s.gbase[digit] + (j - s.tile_base[digit]);
        keys_out[out] = k;
        vals_out[out] = s.vals[j];
    }
}
static_assert(kThreads == kSortBins, "one thread per digit value in the scatter's scan step");

// ------------------------------------------------------------------------------------------------
// u32 exclusive scan (in place), 3 phases, recursive on the tile sums
// ------------------------------------------------------------------------------------------------
constexpr int kScanItems = 4;
constexpr int kScanTile = kThreads * kScanItems;   // 1024

__global__ void __launch_bounds__(kThreads)
k_scan_u32_tile(uint32_t* __restrict__ data, uint32_t n, uint32_t* __restrict__ tile_sums) {
    __shared__ uint32_t warp_tot[kThreads / 32];
    const uint32_t base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    uint32_t v[kScanItems], tot = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        v[k] = base + k < n ? data[base + k] : 0;
        const uint32_t t = v[k];
        v[k] = tot;
        tot += t;
    }
    uint32_t inc = tot;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += y;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    uint32_t woff = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) {
        if (w < warp) woff += warp_tot[w];
        total += warp_tot[w];
    }
    const uint32_t excl = woff + inc - tot;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k)
        if (base + k < n) data[base + k] = v[k] + excl;
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(kThreads)
k_scan_u32_add(uint32_t* __restrict__ data, uint32_t n, const uint32_t* __restrict__ tile_offsets) {
    const uint32_t off = __ldg(tile_offsets + blockIdx.x);
    const uint32_t base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k)
        if (base + k < n) data[base + k] += off;
}

size_t scan_u32_scratch_elems(uint32_t n) {
    size_t total = 0;
    while (n > 1) {
        n = (n + kScanTile - 1) / kScanTile;
        total += n;
        if (n == 1) break;
    }
    return total + 1;
}

// exclusive scan of data[0..n) in place; scratch has scan_u32_scratch_elems(n) elements
int scan_u32(igs_engine* e, uint32_t* data, uint32_t n, uint32_t* scratch, cudaStream_t st) {
    if (n == 0) return IGS_OK;
    const uint32_t nb = (n + kScanTile - 1) / kScanTile;
    k_scan_u32_tile<<<nb, kThreads, 0, st>>>(data, n, scratch);
    IGS_LAUNCHED(e);
    if (nb > 1) {
        int rc = scan_u32(e, scratch, nb, scratch + nb, st);
        if (rc != IGS_OK) return rc;
        k_scan_u32_add<<<nb, kThreads, 0, st>>>(data, n, scratch);
        IGS_LAUNCHED(e);
    }
    return IGS_OK;
}

// Scratch of one job (an index build, a linked-accounts stage), served from a grow-only pool in the engine:
// a job makes the same sequence of requests every time it runs, so a rerun (the rebuild after
// igs_append_events) reuses the buffers of the previous one instead of paying a dozen cudaMalloc / cudaFree
// round trips.
struct Scratch {
    std::vector<igs_buf>& pool;
    size_t next = 0;
    explicit Scratch(std::vector<igs_buf>& p) : pool(p) {}
    template <typename T> cudaError_t alloc(T** p, size_t count) {
        const size_t bytes = (count ? count : 1) * sizeof(T);
        if (next == pool.size()) pool.emplace_back();
        igs_buf& b = pool[next++];
        if (b.cap < bytes) {
            if (b.p) cudaFree(b.p);
            b.p = nullptr; b.cap = 0;
            const size_t cap = ((bytes + bytes / 4) + 255) & ~(size_t)255;
            cudaError_t st = cudaMalloc(&b.p, cap);
            if (st != cudaSuccess) return st;
            b.cap = cap;
        }
        *p = static_cast<T*>(b.p);
        return cudaSuccess;
    }
};

}  // namespace

// stable LSD radix sort of (key, value) pairs on the low passes * digit_bits key bits; returns the index of
// the buffer pair that holds the result
static int radix_sort_pairs(igs_engine* e, uint64_t* keys[2], uint32_t* vals[2], uint32_t n, int passes,
                            int digit_bits, uint32_t* d_hist, uint32_t* d_scan_scratch, cudaStream_t st, int* cur_out) {
    if (digit_bits > kSortMaxBits) {
        igs_set_error(e, "internal: radix digits are at most %d bits", kSortMaxBits);
        return IGS_ERR_INVALID;
    }
    const uint32_t bins = 1u << digit_bits;
    const uint32_t nb_sort = (n + kSortTile - 1) / kSortTile;
    const uint32_t hist_elems = bins * nb_sort;
    // the attribute belongs to the (function, device) pair and the engine's device is current: set it per sort
    // (a few hundred ns) instead of remembering it per process
    IGS_CUDA(e, cudaFuncSetAttribute(k_radix_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SortSmem)));
    int cur = 0;
    for (int p = 0; p < passes; ++p) {
        const int shift = p * digit_bits;
        k_radix_hist<<<nb_sort, kThreads, 0, st>>>(keys[cur], n, shift, bins, d_hist, nb_sort);
        IGS_LAUNCHED(e);
        int rc = scan_u32(e, d_hist, hist_elems, d_scan_scratch, st);
        if (rc != IGS_OK) return rc;
        k_radix_scatter<<<nb_sort, kThreads, sizeof(SortSmem), st>>>(
            keys[cur], vals[cur], keys[cur ^ 1], vals[cur ^ 1], n, shift, bins, d_hist, nb_sort);
        IGS_LAUNCHED(e);
        cur ^= 1;
    }
    *cur_out = cur;
    return IGS_OK;
}

