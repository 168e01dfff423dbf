// k_partition.cu — shard partition of a transaction batch, sm_100a.
//
// shard = splitmix64(account) % n_shards — the key the reference's feature store is addressed by is
// the account (features/redis_store.go:61); every feature, rule and score depends on one account
// only, so shards are independent and no collective follows.  Three kernels, all streaming:
//   1. per-block shard histogram             -> counts[shard][block]       (scratch)
//   2. exclusive scan of counts, shard-major -> bases + offsets[n_shards+1]
//   3. stable scatter of the 32-byte records (+ their original index) to their shard's range.
// Stability keeps the arrival order inside a shard, which the velocity semantics depend on.
#include "igs_internal.cuh"

namespace {

constexpr int kThreads = 512;
constexpr int kItems = 4;
constexpr int kTile = kThreads * kItems;   // 2048 records per block
constexpr int kWarps = kThreads / 32;
constexpr unsigned kMaxShards = 256;

__device__ __forceinline__ uint32_t shard_of(uint32_t account, uint32_t n_shards) {
    uint64_t x = (uint64_t)account + 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    x ^= x >> 31;
    return (uint32_t)(x % n_shards);
}

// one 32-byte record per instruction (LDG.256 / STG.256): a warp reads 1 KB in one request and every store fills its sector
struct Rec32 { uint32_t w[8]; };
static_assert(kMaxShards <= kThreads, "one thread per shard fetches the block's bases");
static_assert(sizeof(igs_tx) == 32 && offsetof(igs_tx, account) == 8, "Rec32.w[2] is igs_tx.account");
__device__ __forceinline__ Rec32 ld_rec32(const void* p) {
    Rec32 r;
    asm("ld.global.nc.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]), "=r"(r.w[7])
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_rec32(void* p, const Rec32& r) {
    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"l"(p), "r"(r.w[0]), "r"(r.w[1]), "r"(r.w[2]), "r"(r.w[3]), "r"(r.w[4]), "r"(r.w[5]), "r"(r.w[6]), "r"(r.w[7])
                 : "memory");
}

__global__ void __launch_bounds__(kThreads)
k_part_hist(const igs_tx* __restrict__ tx, uint32_t n, uint32_t n_shards, uint32_t* __restrict__ counts,
            uint32_t nblocks) {
    __shared__ uint32_t hist[kMaxShards];
    for (uint32_t s = threadIdx.x; s < n_shards; s += kThreads) hist[s] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * kTile;
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
        const uint32_t i = base + r * kThreads + threadIdx.x;
        // one atomic per distinct shard of the warp, not one per lane: with 8 shards 256 lanes hit 8 words
        const uint32_t sh = i < n ? shard_of(__ldg(&tx[i].account), n_shards) : kMaxShards;
        const uint32_t peers = __match_any_sync(0xffffffffu, sh);
        if (i < n && (peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0) atomicAdd(&hist[sh], (uint32_t)__popc(peers));
    }
    __syncthreads();
    for (uint32_t s = threadIdx.x; s < n_shards; s += kThreads) counts[(size_t)s * nblocks + blockIdx.x] = hist[s];
}

// single block: exclusive scan of counts[0..m) in place (m = n_shards * nblocks), shard offsets out.
// 1024 threads, 16 384 counters per round (~50 000 counters — 8 shards of a 12.5M-tx slice — are three rounds).  A warp
// owns 512 consecutive counters and reads them as 16 coalesced rows of 32, carrying its running total from row to row.
constexpr int kScanThreads = 1024, kScanRows = 16, kScanWarps = kScanThreads / 32;
__global__ void __launch_bounds__(kScanThreads)
k_part_scan(uint32_t* __restrict__ counts, uint32_t m, uint32_t n_shards, uint32_t nblocks, uint32_t n,
            uint32_t* __restrict__ offsets) {
    __shared__ uint32_t warp_tot[kScanWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t carry = 0;                             // total of all earlier rounds (every thread keeps its own copy)
    for (uint32_t base = 0; base < m; base += kScanThreads * kScanRows) {
        const uint32_t w0 = base + warp * (32 * kScanRows);
        uint32_t v[kScanRows];
#pragma unroll
        for (int k = 0; k < kScanRows; ++k) {
            const uint32_t i = w0 + k * 32 + lane;
            v[k] = i < m ? counts[i] : 0;
        }
        uint32_t run = 0;                           // exclusive prefix inside the warp's 512 counters
#pragma unroll
        for (int k = 0; k < kScanRows; ++k) {
            uint32_t inc = v[k];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += y;
            }
            const uint32_t row_total = __shfl_sync(0xffffffffu, inc, 31);
            v[k] = run + inc - v[k];
            run += row_total;
        }
        if (lane == 0) warp_tot[warp] = run;
        __syncthreads();
        const uint32_t wt = warp_tot[lane];         // kScanWarps == 32: every warp scans the warp totals itself
        uint32_t winc = wt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += y;
        }
        const uint32_t total = __shfl_sync(0xffffffffu, winc, 31);
        const uint32_t woff = __shfl_sync(0xffffffffu, winc - wt, warp);
#pragma unroll
        for (int k = 0; k < kScanRows; ++k) {
            const uint32_t i = w0 + k * 32 + lane;
            if (i < m) counts[i] = v[k] + carry + woff;
        }
        carry += total;
        __syncthreads();                            // warp_tot is rewritten next round
    }
    __syncthreads();
    for (uint32_t s = threadIdx.x; s < n_shards; s += kScanThreads) offsets[s] = counts[(size_t)s * nblocks];
    if (threadIdx.x == 0) offsets[n_shards] = n;
}
static_assert(kScanWarps == 32, "the warp totals are scanned by one warp-wide shuffle scan");

// kMode 0: 16-byte moves (record arrays not 32-byte aligned).  1: 256-bit moves, records held in registers between
// the ranking and the store.  (Reading the records a second time for the store instead — a third of the registers,
// twice the resident warps — measured 3 % slower: the kernel is not short of warps, it runs at ~76 % of the copy
// bandwidth on one read stream and 2 x n_shards write streams.)
template <int kMode>
__global__ void __launch_bounds__(kThreads, 2)
k_part_scatter(const igs_tx* __restrict__ tx, uint32_t n, uint32_t n_shards,
               const uint32_t* __restrict__ bases, uint32_t nblocks, igs_tx* __restrict__ tx_out,
               uint32_t* __restrict__ perm) {
    __shared__ uint32_t cnt[kWarps * kMaxShards];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (uint32_t d = threadIdx.x; d < n_shards * kWarps; d += kThreads) cnt[d] = 0;
    __syncthreads();
    uint32_t* my_cnt = cnt + warp * n_shards;
    const uint32_t wbase = blockIdx.x * kTile + warp * (kItems * 32);
    constexpr bool kWide = kMode == 1;
    uint32_t shard[kItems], rank[kItems];
    Rec32 rec[kWide ? kItems : 1];
    // this block's start in each shard's range: asked for now, needed after the ranking
    uint32_t my_base = 0;
    if (threadIdx.x < n_shards) my_base = __ldg(bases + (size_t)threadIdx.x * nblocks + blockIdx.x);
    const uint32_t lt_mask = (1u << lane) - 1u;
    if (kWide) {                                   // all eight loads in flight before the first ranking round needs one
#pragma unroll
        for (int r = 0; r < kItems; ++r) {
            const uint32_t i = wbase + r * 32 + lane;
            rec[r] = ld_rec32(tx + (i < n ? i : n - 1));
        }
    }
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
        const uint32_t i = wbase + r * 32 + lane;
        const bool valid = i < n;
        const uint32_t account = kWide ? rec[r].w[2] /* igs_tx.account: bytes 8..11 */ : valid ? __ldg(&tx[i].account) : 0u;
        shard[r] = valid ? shard_of(account, n_shards) : (kMaxShards + lane);
        const uint32_t peers = __match_any_sync(0xffffffffu, shard[r]);
        const uint32_t before = valid ? my_cnt[shard[r]] : 0;
        __syncwarp();
        if (valid && (peers & lt_mask) == 0) my_cnt[shard[r]] = before + __popc(peers);
        __syncwarp();
        rank[r] = before + __popc(peers & lt_mask);
    }
    __syncthreads();
    for (uint32_t s = threadIdx.x; s < n_shards; s += kThreads) {      // (n_shards <= kMaxShards <= kThreads: one round)
        uint32_t run = my_base;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
            const uint32_t c = cnt[w * n_shards + s];
            cnt[w * n_shards + s] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
        const uint32_t i = wbase + r * 32 + lane;
        if (i < n) {
            const uint32_t pos = my_cnt[shard[r]] + rank[r];
            if (kWide) {
                st_rec32(tx_out + pos, rec[r]);
            } else {
                const uint4* src = reinterpret_cast<const uint4*>(tx + i);
                uint4* dst = reinterpret_cast<uint4*>(tx_out + pos);
                dst[0] = __ldg(src + 0);
                dst[1] = __ldg(src + 1);
            }
            if (perm) perm[pos] = i;
        }
    }
}

}  // namespace

// Runs the three kernels with `counts` (>= n_shards * ceil(n / 2048) words) as their scratch.
int igs_partition_run(igs_engine* e, const igs_tx* d_tx, uint64_t n64, uint32_t n_shards, uint32_t* counts,
                      igs_tx* d_tx_out, uint32_t* d_perm, uint32_t* d_offsets, cudaStream_t stream) {
    const uint32_t n = (uint32_t)n64;
    const uint32_t nblocks = (uint32_t)((n64 + kTile - 1) / kTile);    // 64-bit: n + kTile - 1 wraps near 2^32
    if (n == 0) {
        IGS_CUDA(e, cudaMemsetAsync(d_offsets, 0, ((size_t)n_shards + 1) * sizeof(uint32_t), stream));
        return IGS_OK;
    }
    k_part_hist<<<nblocks, kThreads, 0, stream>>>(d_tx, n, n_shards, counts, nblocks);
    IGS_LAUNCHED(e);
    k_part_scan<<<1, kScanThreads, 0, stream>>>(counts, n_shards * nblocks, n_shards, nblocks, n, d_offsets);
    IGS_LAUNCHED(e);
    if ((((uintptr_t)d_tx | (uintptr_t)d_tx_out) & 31u) != 0)
        k_part_scatter<0><<<nblocks, kThreads, 0, stream>>>(d_tx, n, n_shards, counts, nblocks, d_tx_out, d_perm);
    else
        k_part_scatter<1><<<nblocks, kThreads, 0, stream>>>(d_tx, n, n_shards, counts, nblocks, d_tx_out, d_perm);
    IGS_LAUNCHED(e);
    return IGS_OK;
}

size_t igs_partition_scratch_words(uint64_t n, uint32_t n_shards) {
    return (size_t)n_shards * ((n + kTile - 1) / kTile);
}

// ------------------------------------------------------------------------------------------------
// igs_partition_tx_static — the same three kernels, with their counters in memory that is part of
// the library image instead of memory obtained from the allocator.
// ------------------------------------------------------------------------------------------------
// Who needs this.  A risk-service shard partitions incoming batches on its ingest thread while
// scoring batches of other slots are in flight on the same GPU.  igs_partition_tx_device takes its
// counters from the engine, which calls cudaMalloc the first time and again whenever a batch needs
// more counters than any batch before it (cudaFree + cudaMalloc).  Both calls serialise against
// everything running on the device, so that one ingest call would stall every in-flight scoring
// call and surface as a latency spike.  The other such caller is a deployment that fixes its GPU
// memory budget at start-up (a shard sharing the GPU with another tenant) and treats any later
// allocation as a fault.  Neither can afford an allocating call on this path.
//
// The counters therefore live in the array below: 1 MiB that every GPU which loads this library
// holds from module load until the process exits, whether or not this entry point is ever used.
#define IGS_PARTITION_STATIC_WORDS_ (256u * 1024u)
static_assert(IGS_PARTITION_STATIC_WORDS_ == IGS_PARTITION_STATIC_WORDS, "keep igs.h in step");
static __device__ __align__(256) uint32_t g_part_counters[IGS_PARTITION_STATIC_WORDS_];

// In host code `g_part_counters` is only a handle for the runtime, not a device address; the
// address has to be asked for, once per device and with that device current.
namespace {
constexpr int kMaxDevices = 64;
uint32_t*  g_part_counters_dev[kMaxDevices] = {nullptr};
std::mutex g_part_mu;
}  // namespace

extern "C" int igs_partition_tx_static(igs_engine* e, const igs_tx* d_tx, uint64_t n, uint32_t n_shards,
                                       igs_tx* d_tx_out, uint32_t* d_perm, uint32_t* d_offsets, void* stream) {
    if (!e) {
        igs_set_error(nullptr, "engine handle is null");
        return IGS_ERR_INVALID;
    }
    IGS_REQUIRE(e, e->device >= 0 && e->device < kMaxDevices, "device ordinal beyond 63");
    IGS_CUDA(e, cudaSetDevice(e->device));
    IGS_REQUIRE(e, n_shards >= 1 && n_shards <= 256, "n_shards must be 1..256");
    IGS_REQUIRE(e, n < (1ull << 32), "at most 2^32-1 records per call");
    IGS_REQUIRE(e, d_offsets != nullptr, "offsets is null");
    IGS_REQUIRE(e, n == 0 || (d_tx && d_tx_out), "tx / tx_out is null");
    const size_t words = igs_partition_scratch_words(n, n_shards);
    if (words > IGS_PARTITION_STATIC_WORDS_) {
        igs_set_error(e,
                      "igs_partition_tx_static: n_shards * ceil(n / 2048) = %zu counters, the built-in array "
                      "holds %u; split the batch or use igs_partition_tx_device",
                      words, IGS_PARTITION_STATIC_WORDS_);
        return IGS_ERR_INVALID;
    }
    uint32_t* counters;
    {
        std::lock_guard<std::mutex> lk(g_part_mu);
        if (!g_part_counters_dev[e->device]) {
            // first call on this device: resolve the array's address there
            uint32_t* p = nullptr;
            IGS_CUDA(e, cudaGetSymbolAddress((void**)&p, g_part_counters));
            g_part_counters_dev[e->device] = p;
        }
        counters = g_part_counters_dev[e->device];
    }
    // Nothing has to be cleared: the histogram kernel writes every counter the scan and the scatter
    // read (n_shards * ceil(n / 2048) words), whatever an earlier call left behind.
    return igs_partition_run(e, d_tx, n, n_shards, counters, d_tx_out, d_perm, d_offsets, (cudaStream_t)stream);
}
