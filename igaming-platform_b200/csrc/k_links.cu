// k_links.cu — linked accounts and the bonus-abuse check, the one step that crosses shards.
//
// Serves RiskChecker.CheckBonusAbuse (services/bonus/internal/service/bonus_engine.go:139-141,268-275) and
// rpc CheckBonusAbuse (proto/risk/v1/risk.proto:20,135-145).  The reference declares the interfaces only; the
// semantics are the spec in include/igs.h.  Accounts are sharded by account, devices are not, so two accounts
// that share a device can sit on different GPUs: the links are found in three device-side stages with an
// all-to-all between them (the caller's collective — NCCL over NVLink in links.py):
//   stage 1  igs_links_pairs   local history -> (device, account) pairs grouped by the device's owner shard
//   stage 2  igs_links_group   owner: distinct pairs, per-device account sets, directed links (account, other)
//                              grouped by the account's shard
//   stage 3  igs_links_finish  account owner: distinct links -> per-account {count, three smallest ids}
// All of it is HBM-streaming integer work: hash-keyed stable radix sorts (40-bit keys, 5 passes) where equal
// hashes are resolved by comparing the full 128-bit records, so the results are exact.
#include <string.h>

#include <algorithm>
#include <vector>

#include "idset.cuh"
#include "igs_internal.cuh"
#include "sortscan.cuh"

namespace {

constexpr int kKeyBits = 40;          // 5 radix passes of 8 bits
constexpr int kKeyPasses = 5;
constexpr int kMaxShards = 256;
// Slots of the engine's links pool: every stage makes the same requests in the same order on every run and
// owns a fixed range, so a rerun reuses its buffers and a stage never overwrites the previous stage's output
// (with one shard the caller hands stage N's output straight to stage N + 1).
constexpr size_t kStage2First = 2;      // stage 1: counters, pairs
constexpr size_t kStage3First = 40;     // stage 2 uses fewer than 38 slots

__host__ __device__ __forceinline__ uint32_t shard_of(uint64_t x, uint32_t n_shards) {   // = igs_shard_of(x, 0, n)
    return (uint32_t)(igs_idset_mix(x + 0x9E3779B97F4A7C15ull) % n_shards);
}
__device__ __forceinline__ uint64_t key40(uint64_t a, uint64_t b) {
    return igs_idset_mix(igs_idset_mix(a) ^ (b * 0x9E3779B97F4A7C15ull)) >> (64 - kKeyBits);
}

// ---- stage 1 ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
k_lp_count(const uint64_t* __restrict__ device, uint32_t n, uint32_t n_shards, unsigned long long* __restrict__ counts) {
    __shared__ uint32_t sh[kMaxShards];
    for (uint32_t s = threadIdx.x; s < n_shards; s += kThreads) sh[s] = 0;
    __syncthreads();
    for (uint32_t i = blockIdx.x * kThreads + threadIdx.x; i < n; i += gridDim.x * kThreads) {
        const uint64_t d = __ldg(device + i);
        if (d) atomicAdd(&sh[shard_of(d, n_shards)], 1u);
    }
    __syncthreads();
    for (uint32_t s = threadIdx.x; s < n_shards; s += kThreads)
        if (sh[s]) atomicAdd(counts + s, (unsigned long long)sh[s]);
}

// One cursor per destination shard; a warp claims the slots of all its lanes that go to the same shard
// with a single atomic.  The order inside a group is arbitrary — the next stage sorts.
__global__ void __launch_bounds__(kThreads)
k_lp_scatter(const uint64_t* __restrict__ device, const uint32_t* __restrict__ account, uint32_t n,
             uint32_t shard, uint32_t n_shards, unsigned long long* __restrict__ cursor,
             igs_link_pair* __restrict__ out) {
    const uint32_t i = blockIdx.x * kThreads + threadIdx.x;
    const uint64_t d = i < n ? __ldg(device + i) : 0;
    const uint32_t dest = d ? shard_of(d, n_shards) : 0xFFFFFFFFu;
    const unsigned peers = __match_any_sync(0xffffffffu, dest);
    if (!d) return;
    const int lane = threadIdx.x & 31, leader = __ffs(peers) - 1;
    unsigned long long base = 0;
    if (lane == leader) base = atomicAdd(cursor + dest, (unsigned long long)__popc(peers));
    base = __shfl_sync(peers, base, leader);
    igs_link_pair p;
    p.device = d;
    p.account = ((uint64_t)shard << 32) | __ldg(account + i);
    out[base + __popc(peers & ((1u << lane) - 1u))] = p;
}

// ---- shared by stages 2 and 3: exact dedupe of 16-byte records through a hash-keyed sort ----------------
struct Rec { uint64_t a, b; };        // igs_link_pair {device, account} or igs_link {account, other}

__global__ void k_rec_keys(const Rec* __restrict__ rec, uint32_t n, int by_first_only, uint64_t* __restrict__ keys,
                           uint32_t* __restrict__ vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Rec r = rec[i];
    keys[i] = by_first_only ? igs_idset_mix(r.a) >> (64 - kKeyBits) : key40(r.a, r.b);
    vals[i] = i;
}
// first[r] = 1 iff no earlier entry of the sorted order holds the same record
__global__ void k_rec_first(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                            const Rec* __restrict__ rec, uint32_t n, uint32_t* __restrict__ first) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const uint64_t key = __ldg(keys + r);
    const Rec me = rec[__ldg(vals + r)];
    uint32_t f = 1;
    for (uint32_t rr = r; rr-- > 0;) {
        if (__ldg(keys + rr) != key) break;
        const Rec o = rec[__ldg(vals + rr)];
        if (o.a == me.a && o.b == me.b) { f = 0; break; }
    }
    first[r] = f;
}
// ---- stage 2 ------------------------------------------------------------------------------------
__global__ void k_gather_rec(const uint32_t* __restrict__ vals, const Rec* __restrict__ rec, uint32_t n,
                             Rec* __restrict__ out) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n) out[r] = rec[__ldg(vals + r)];
}
__global__ void k_compact_flagged(const uint32_t* __restrict__ vals, const Rec* __restrict__ rec,
                                  const uint32_t* __restrict__ flag_ex, const uint32_t* __restrict__ last_flag,
                                  uint32_t n, Rec* __restrict__ out) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const uint32_t ex = __ldg(flag_ex + r);
    const bool flagged = r + 1 < n ? __ldg(flag_ex + r + 1) != ex : *last_flag != 0;
    if (flagged) out[ex] = rec[__ldg(vals + r)];
}
__global__ void k_save_last(const uint32_t* __restrict__ flag, uint32_t n, uint32_t* __restrict__ last_flag) {
    if (blockIdx.x == 0 && threadIdx.x == 0) *last_flag = flag[n - 1];
}

// pairs sorted by the hash of their device (keys) — accounts of one device share a run.  links_of[r] = number
// of links pair r emits: k - 1 when its device has 2 <= k <= IGS_LINK_MAX_ACCOUNTS accounts, else 0.  The walk
// stops as soon as the device is known to be public, so a kiosk device costs 65 steps per pair, not k.
__device__ __forceinline__ uint32_t accounts_of_device(const uint64_t* __restrict__ keys, const Rec* __restrict__ pairs,
                                                       uint32_t n, uint32_t r) {
    const uint64_t key = __ldg(keys + r), dev = pairs[r].a;
    uint32_t k = 1;
    for (uint32_t rr = r; rr-- > 0 && k <= IGS_LINK_MAX_ACCOUNTS;) {
        if (__ldg(keys + rr) != key) break;
        k += pairs[rr].a == dev;
    }
    for (uint32_t rr = r + 1; rr < n && k <= IGS_LINK_MAX_ACCOUNTS; ++rr) {
        if (__ldg(keys + rr) != key) break;
        k += pairs[rr].a == dev;
    }
    return k;
}
__global__ void k_lg_count(const uint64_t* __restrict__ keys, const Rec* __restrict__ pairs, uint32_t n,
                           uint32_t n_shards, uint32_t* __restrict__ links_of, unsigned long long* __restrict__ counts) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const uint32_t k = accounts_of_device(keys, pairs, n, r);
    const uint32_t out = (k >= 2 && k <= IGS_LINK_MAX_ACCOUNTS) ? k - 1 : 0;
    links_of[r] = out;
    if (out) atomicAdd(counts + min((uint32_t)(pairs[r].b >> 32), n_shards - 1), (unsigned long long)out);
}
__global__ void k_lg_emit(const uint64_t* __restrict__ keys, const Rec* __restrict__ pairs, uint32_t n,
                          uint32_t n_shards, const uint32_t* __restrict__ links_of,
                          unsigned long long* __restrict__ cursor, igs_link* __restrict__ out) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const uint32_t cnt = __ldg(links_of + r);
    if (!cnt) return;
    const Rec me = pairs[r];
    unsigned long long pos = atomicAdd(cursor + min((uint32_t)(me.b >> 32), n_shards - 1), (unsigned long long)cnt);
    const uint64_t key = __ldg(keys + r);
    for (uint32_t rr = r; rr-- > 0;) {
        if (__ldg(keys + rr) != key) break;
        const Rec o = pairs[rr];
        if (o.a == me.a) { igs_link l; l.account = me.b; l.other = o.b; out[pos++] = l; }
    }
    for (uint32_t rr = r + 1; rr < n; ++rr) {
        if (__ldg(keys + rr) != key) break;
        const Rec o = pairs[rr];
        if (o.a == me.a) { igs_link l; l.account = me.b; l.other = o.b; out[pos++] = l; }
    }
}

// ---- peer-memory variants: the exchange is part of the producing kernel ----------------------------------
// A shard's receive window as its peers see it.  cursors[0] / cursors[1] count the pair / link slots claimed
// so far (they keep counting past the capacity, so an overflow reports the size that was needed).
struct PeerWin {
    unsigned long long* cursors;
    Rec* pairs;
    Rec* links;
};
constexpr size_t kWindowHeader = 256;
__host__ PeerWin window_view(void* base, uint64_t cap_pairs) {
    PeerWin w;
    w.cursors = static_cast<unsigned long long*>(base);
    w.pairs = reinterpret_cast<Rec*>(static_cast<char*>(base) + kWindowHeader);
    w.links = w.pairs + cap_pairs;
    return w;
}

// stage 1 + exchange 1: each pair goes straight into the window of the shard that owns its device — over
// NVLink when that shard is another GPU.  A block first ranks its 1024 events per destination in shared
// memory, then claims each destination's slots with ONE system-scope atomic, so the remote cursors see a
// few atomics per thousand records.
constexpr int kPeerItems = 4;
__global__ void __launch_bounds__(kThreads)
k_lp_scatter_peer(const uint64_t* __restrict__ device, const uint32_t* __restrict__ account, uint32_t n,
                  uint32_t shard, uint32_t n_shards, const PeerWin* __restrict__ wins, unsigned long long cap_pairs) {
    __shared__ uint32_t sh_cnt[kMaxShards];
    __shared__ unsigned long long sh_base[kMaxShards];
    for (uint32_t s = threadIdx.x; s < n_shards; s += kThreads) sh_cnt[s] = 0;
    __syncthreads();
    uint64_t d[kPeerItems];
    uint32_t dest[kPeerItems], rank[kPeerItems];
    const uint32_t base_i = blockIdx.x * (kThreads * kPeerItems) + threadIdx.x;
#pragma unroll
    for (int k = 0; k < kPeerItems; ++k) {
        const uint32_t i = base_i + k * kThreads;
        d[k] = i < n ? __ldg(device + i) : 0;
        dest[k] = d[k] ? shard_of(d[k], n_shards) : 0;
        rank[k] = d[k] ? atomicAdd(&sh_cnt[dest[k]], 1u) : 0;
    }
    __syncthreads();
    for (uint32_t s = threadIdx.x; s < n_shards; s += kThreads)
        sh_base[s] = sh_cnt[s] ? atomicAdd_system(wins[s].cursors + 0, (unsigned long long)sh_cnt[s]) : 0ull;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kPeerItems; ++k) {
        if (!d[k]) continue;
        const unsigned long long slot = sh_base[dest[k]] + rank[k];
        if (slot < cap_pairs) {
            Rec p;
            p.a = d[k];
            p.b = ((uint64_t)shard << 32) | __ldg(account + base_i + k * kThreads);
            wins[dest[k]].pairs[slot] = p;
        }
    }
}

// stage 2's emit + exchange 2: the links of a pair go straight into the window of the account's shard; the
// block sums its link counts per destination and claims them with one system-scope atomic each
__global__ void __launch_bounds__(kThreads)
k_lg_emit_peer(const uint64_t* __restrict__ keys, const Rec* __restrict__ pairs, uint32_t n,
               uint32_t n_shards, const uint32_t* __restrict__ links_of,
               const PeerWin* __restrict__ wins, unsigned long long cap_links) {
    __shared__ uint32_t sh_cnt[kMaxShards];
    __shared__ unsigned long long sh_base[kMaxShards];
    for (uint32_t s = threadIdx.x; s < n_shards; s += kThreads) sh_cnt[s] = 0;
    __syncthreads();
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t cnt = r < n ? __ldg(links_of + r) : 0;
    Rec me;
    me.a = me.b = 0;
    uint32_t dst = 0, off = 0;
    if (cnt) {
        me = pairs[r];
        dst = min((uint32_t)(me.b >> 32), n_shards - 1);
        off = atomicAdd(&sh_cnt[dst], cnt);
    }
    __syncthreads();
    for (uint32_t s = threadIdx.x; s < n_shards; s += kThreads)
        sh_base[s] = sh_cnt[s] ? atomicAdd_system(wins[s].cursors + 1, (unsigned long long)sh_cnt[s]) : 0ull;
    __syncthreads();
    if (!cnt) return;
    unsigned long long pos = sh_base[dst] + off;
    if (pos + cnt > cap_links) return;                        // the consumer reports the overflow
    Rec* out = wins[dst].links;
    const uint64_t key = __ldg(keys + r);
    for (uint32_t rr = r; rr-- > 0;) {
        if (__ldg(keys + rr) != key) break;
        const Rec o = pairs[rr];
        if (o.a == me.a) { Rec l; l.a = me.b; l.b = o.b; out[pos++] = l; }
    }
    for (uint32_t rr = r + 1; rr < n; ++rr) {
        if (__ldg(keys + rr) != key) break;
        const Rec o = pairs[rr];
        if (o.a == me.a) { Rec l; l.a = me.b; l.b = o.b; out[pos++] = l; }
    }
}

// ---- stage 3 ------------------------------------------------------------------------------------
__global__ void k_lf_init(igs_links_row* __restrict__ rows, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    igs_links_row r;
    r.n_linked = 0; r.reserved = 0;
    r.linked[0] = r.linked[1] = r.linked[2] = ~0ull;
    rows[i] = r;
}
// round 0 counts and finds the smallest id; rounds 1, 2 the next ones (ids above the previous minimum)
__global__ void k_lf_round(const Rec* __restrict__ links, uint32_t n, uint32_t shard, uint32_t n_rows, int round,
                           igs_links_row* __restrict__ rows) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Rec l = links[i];
    const uint32_t idx = (uint32_t)l.a;
    if ((uint32_t)(l.a >> 32) != shard || idx >= n_rows) return;      // not an account of this shard
    igs_links_row* row = rows + idx;
    if (round == 0) atomicAdd(&row->n_linked, 1u);
    if (round == 0 || l.b > row->linked[round - 1])
        atomicMin(reinterpret_cast<unsigned long long*>(&row->linked[round]), (unsigned long long)l.b);
}
__global__ void k_lf_close(igs_links_row* __restrict__ rows, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
#pragma unroll
    for (int k = 0; k < 3; ++k)
        if (rows[i].linked[k] == ~0ull) rows[i].linked[k] = 0;
}

// ---- queries ------------------------------------------------------------------------------------
__global__ void k_links_get(const igs_links_row* __restrict__ rows, uint64_t n_rows, const uint32_t* __restrict__ acct,
                            uint64_t n, igs_links_row* __restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t a = __ldg(acct + i);
    igs_links_row r;
    r.n_linked = 0; r.reserved = 0; r.linked[0] = r.linked[1] = r.linked[2] = 0;
    if (a < n_rows) r = rows[a];
    out[i] = r;
}
__global__ void k_bonus_abuse(const igs_links_row* __restrict__ rows, uint64_t n_rows,
                              const igs_player_row* __restrict__ players, uint64_t n_players,
                              const uint32_t* __restrict__ acct, uint64_t n, igs_abuse* __restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t a = __ldg(acct + i);
    bool bonus_only = false;
    if (a < n_players)                                                   // scoring/engine.go:384-386
        bonus_only = players[a].bonus_claim_count > 3 && players[a].total_deposits < 5000;
    const uint32_t nl = a < n_rows ? rows[a].n_linked : 0;
    const int tenths = 5 * (int)bonus_only + 3 * (int)(nl >= 1) + 2 * (int)(nl >= 3);
    igs_abuse r;
    r.abuse_score = __fdiv_rn((float)tenths, 10.0f);
    r.n_linked = nl;
    r.is_abuser = tenths >= 5;
    r.signals = (uint8_t)((bonus_only ? IGS_ABUSE_BONUS_ONLY : 0u) | (nl >= 1 ? IGS_ABUSE_MULTI_ACCOUNT : 0u) |
                          (nl >= 3 ? IGS_ABUSE_ACCOUNT_RING : 0u));
#pragma unroll
    for (int k = 0; k < 6; ++k) r.pad[k] = 0;
    out[i] = r;
}

// group sizes -> group starts; returns the total
uint64_t offsets_of(const unsigned long long* counts, uint32_t n_shards, unsigned long long* starts) {
    uint64_t run = 0;
    for (uint32_t s = 0; s < n_shards; ++s) { starts[s] = run; run += counts[s]; }
    return run;
}

// Sorts n 16-byte records by a 40-bit hash (of the whole record, or of its first word only), leaving the
// sorted keys / permutation in keys[cur] / vals[cur].
int sort_records(igs_engine* e, Scratch& sc, const Rec* d_rec, uint32_t n, int by_first_only, uint64_t* keys[2],
                 uint32_t* vals[2], int* cur, cudaStream_t st) {
    const uint32_t nb_sort = (n + kSortTile - 1) / kSortTile;
    uint32_t *d_hist, *d_scr;
    IGS_CUDA(e, sc.alloc(&keys[0], n));
    IGS_CUDA(e, sc.alloc(&keys[1], n));
    IGS_CUDA(e, sc.alloc(&vals[0], n));
    IGS_CUDA(e, sc.alloc(&vals[1], n));
    IGS_CUDA(e, sc.alloc(&d_hist, (size_t)kSortBins * nb_sort));
    IGS_CUDA(e, sc.alloc(&d_scr, scan_u32_scratch_elems(std::max<uint32_t>((uint32_t)kSortBins * nb_sort, n + 1))));
    *cur = 0;
    if (n == 0) return IGS_OK;
    k_rec_keys<<<(n + kThreads - 1) / kThreads, kThreads, 0, st>>>(d_rec, n, by_first_only, keys[0], vals[0]);
    IGS_LAUNCHED(e);
    return radix_sort_pairs(e, keys, vals, n, kKeyPasses, kKeyBits / kKeyPasses, d_hist, d_scr, st, cur);
}

// Distinct records of d_rec[0..n) -> d_out (sc-owned), *n_out.  Exact.
int distinct_records(igs_engine* e, Scratch& sc, const Rec* d_rec, uint32_t n, Rec** d_out, uint32_t* n_out,
                     cudaStream_t st) {
    uint64_t* keys[2]; uint32_t* vals[2];
    int cur = 0;
    int rc = sort_records(e, sc, d_rec, n, 0, keys, vals, &cur, st);
    if (rc != IGS_OK) return rc;
    uint32_t *first, *last_flag, *scr;
    IGS_CUDA(e, sc.alloc(&first, n));
    IGS_CUDA(e, sc.alloc(&last_flag, 1));
    IGS_CUDA(e, sc.alloc(&scr, scan_u32_scratch_elems(n + 1)));
    IGS_CUDA(e, sc.alloc(d_out, n));
    *n_out = 0;
    if (n == 0) return IGS_OK;
    const uint32_t nb = (n + kThreads - 1) / kThreads;
    k_rec_first<<<nb, kThreads, 0, st>>>(keys[cur], vals[cur], d_rec, n, first);
    IGS_LAUNCHED(e);
    k_save_last<<<1, 32, 0, st>>>(first, n, last_flag);
    rc = scan_u32(e, first, n, scr, st);
    if (rc != IGS_OK) return rc;
    k_compact_flagged<<<nb, kThreads, 0, st>>>(vals[cur], d_rec, first, last_flag, n, *d_out);
    IGS_LAUNCHED(e);
    uint32_t ex = 0, lf = 0;
    IGS_CUDA(e, cudaMemcpyAsync(&ex, first + (n - 1), 4, cudaMemcpyDeviceToHost, st));
    IGS_CUDA(e, cudaMemcpyAsync(&lf, last_flag, 4, cudaMemcpyDeviceToHost, st));
    IGS_CUDA(e, cudaStreamSynchronize(st));
    *n_out = ex + lf;
    return IGS_OK;
}

}  // namespace

void igs_links_free(igs_engine* e) {
    for (igs_buf& b : e->links_pool)
        if (b.p) cudaFree(b.p);
    e->links_pool.clear();
    if (e->d_links_rows) cudaFree(e->d_links_rows);
    e->d_links_rows = nullptr;
    e->n_links_rows = 0;
    for (int s = 0; s < kMaxShards; ++s) {
        if (e->links_peer[s] && e->links_peer_ipc[s]) cudaIpcCloseMemHandle(e->links_peer[s]);
        e->links_peer[s] = nullptr;
        e->links_peer_ipc[s] = false;
    }
    if (e->links_window) cudaFree(e->links_window);
    e->links_window = nullptr;
}

extern "C" {

int igs_links_pairs(igs_engine* e, uint32_t shard, uint32_t n_shards, const igs_link_pair** d_pairs, uint64_t* counts) {
    if (!e) return IGS_ERR_INVALID;
    IGS_REQUIRE(e, d_pairs && counts, "d_pairs / counts is null");
    IGS_REQUIRE(e, n_shards >= 1 && n_shards <= kMaxShards && shard < n_shards, "shard / n_shards out of range");
    if (!e->has_events || !e->has_event_ids) {
        igs_set_error(e, "no event history with identifiers loaded (igs_load_events_ids)");
        return IGS_ERR_STATE;
    }
    IGS_CUDA(e, cudaSetDevice(e->device));
    std::lock_guard<std::mutex> lk(e->mu);
    cudaStream_t st = e->slot[0].stream;
    const uint32_t n = (uint32_t)e->n_raw;
    Scratch sc(e->links_pool);
    unsigned long long* d_counts;
    igs_link_pair* d_out;
    IGS_CUDA(e, sc.alloc(&d_counts, kMaxShards));
    IGS_CUDA(e, sc.alloc(&d_out, n));
    IGS_CUDA(e, cudaMemsetAsync(d_counts, 0, kMaxShards * 8, st));
    unsigned long long h_counts[kMaxShards] = {0}, h_starts[kMaxShards];
    if (n) {
        const uint32_t nb = (n + kThreads - 1) / kThreads;
        k_lp_count<<<std::min<uint32_t>(nb, (uint32_t)e->sm_count * 8), kThreads, 0, st>>>(e->raw_device, n, n_shards, d_counts);
        IGS_LAUNCHED(e);
        IGS_CUDA(e, cudaMemcpyAsync(h_counts, d_counts, n_shards * 8, cudaMemcpyDeviceToHost, st));
        IGS_CUDA(e, cudaStreamSynchronize(st));
        offsets_of(h_counts, n_shards, h_starts);
        IGS_CUDA(e, cudaMemcpyAsync(d_counts, h_starts, n_shards * 8, cudaMemcpyHostToDevice, st));
        k_lp_scatter<<<nb, kThreads, 0, st>>>(e->raw_device, e->raw_account, n, shard, n_shards, d_counts, d_out);
        IGS_LAUNCHED(e);
        IGS_CUDA(e, cudaStreamSynchronize(st));
    }
    for (uint32_t s = 0; s < n_shards; ++s) counts[s] = h_counts[s];
    *d_pairs = d_out;
    return IGS_OK;
}

// Stage 2.  d_wins == nullptr: links into an engine-owned buffer grouped by shard (*d_links, counts);
// else: links straight into the account owners' windows.
static int links_group_impl(igs_engine* e, const igs_link_pair* d_pairs, uint64_t n64, uint32_t n_shards,
                            const igs_link** d_links, uint64_t* counts, const PeerWin* d_wins) {
    IGS_REQUIRE(e, n64 == 0 || d_pairs, "d_pairs is null");
    IGS_REQUIRE(e, n_shards >= 1 && n_shards <= kMaxShards, "n_shards out of range");
    IGS_REQUIRE(e, n64 < 0xFFFFFFFFull, "at most 2^32-2 pairs per stage");
    cudaStream_t st = e->slot[0].stream;
    const uint32_t n = (uint32_t)n64;
    Scratch sc(e->links_pool);
    if (e->links_pool.size() < kStage2First) e->links_pool.resize(kStage2First);
    sc.next = kStage2First;                             // keep stage 1's output alive (one shard: d_pairs is it)
    Rec* d_distinct; uint32_t m = 0;
    int rc = distinct_records(e, sc, reinterpret_cast<const Rec*>(d_pairs), n, &d_distinct, &m, st);
    if (rc != IGS_OK) return rc;
    // order the distinct pairs by device hash and make them contiguous in that order
    uint64_t* keys[2]; uint32_t* vals[2];
    int cur = 0;
    rc = sort_records(e, sc, d_distinct, m, 1, keys, vals, &cur, st);
    if (rc != IGS_OK) return rc;
    Rec* d_sorted; uint32_t* d_links_of; unsigned long long* d_counts;
    IGS_CUDA(e, sc.alloc(&d_sorted, m));
    IGS_CUDA(e, sc.alloc(&d_links_of, m));
    IGS_CUDA(e, sc.alloc(&d_counts, kMaxShards));
    IGS_CUDA(e, cudaMemsetAsync(d_counts, 0, kMaxShards * 8, st));
    unsigned long long h_counts[kMaxShards] = {0}, h_starts[kMaxShards];
    uint64_t total = 0;
    const uint32_t mb = (m + kThreads - 1) / kThreads;
    if (m) {
        k_gather_rec<<<mb, kThreads, 0, st>>>(vals[cur], d_distinct, m, d_sorted);
        IGS_LAUNCHED(e);
        k_lg_count<<<mb, kThreads, 0, st>>>(keys[cur], d_sorted, m, n_shards, d_links_of, d_counts);
        IGS_LAUNCHED(e);
        if (d_wins) {
            k_lg_emit_peer<<<mb, kThreads, 0, st>>>(keys[cur], d_sorted, m, n_shards, d_links_of, d_wins,
                                                    (unsigned long long)e->links_cap_links);
            IGS_LAUNCHED(e);
        }
        IGS_CUDA(e, cudaMemcpyAsync(h_counts, d_counts, n_shards * 8, cudaMemcpyDeviceToHost, st));
        IGS_CUDA(e, cudaStreamSynchronize(st));
        total = offsets_of(h_counts, n_shards, h_starts);
    }
    if (counts)
        for (uint32_t s = 0; s < n_shards; ++s) counts[s] = h_counts[s];
    if (d_wins) return IGS_OK;
    igs_link* d_out;
    IGS_CUDA(e, sc.alloc(&d_out, total));
    if (sc.next > kStage3First) {
        igs_set_error(e, "internal: stage 2 outgrew its scratch slots");
        return IGS_ERR_STATE;
    }
    if (total) {
        IGS_CUDA(e, cudaMemcpyAsync(d_counts, h_starts, n_shards * 8, cudaMemcpyHostToDevice, st));
        k_lg_emit<<<mb, kThreads, 0, st>>>(keys[cur], d_sorted, m, n_shards, d_links_of, d_counts, d_out);
        IGS_LAUNCHED(e);
        IGS_CUDA(e, cudaStreamSynchronize(st));
    }
    *d_links = d_out;
    return IGS_OK;
}

int igs_links_group(igs_engine* e, const igs_link_pair* d_pairs, uint64_t n64, uint32_t n_shards,
                    const igs_link** d_links, uint64_t* counts) {
    if (!e) return IGS_ERR_INVALID;
    IGS_REQUIRE(e, d_links && counts, "d_links / counts is null");
    IGS_CUDA(e, cudaSetDevice(e->device));
    std::lock_guard<std::mutex> lk(e->mu);
    return links_group_impl(e, d_pairs, n64, n_shards, d_links, counts, nullptr);
}

static int links_finish_impl(igs_engine* e, const igs_link* d_links, uint64_t n64, uint32_t shard) {
    IGS_REQUIRE(e, n64 == 0 || d_links, "d_links is null");
    IGS_REQUIRE(e, n64 < 0xFFFFFFFFull, "at most 2^32-2 links per stage");
    if (!e->has_events) {
        igs_set_error(e, "no event history loaded: the account index space is unknown (igs_load_events_ids)");
        return IGS_ERR_STATE;
    }
    cudaStream_t st = e->slot[0].stream;
    const uint32_t n = (uint32_t)n64;
    const uint32_t n_rows = e->ev.n_accounts;
    IGS_CUDA(e, cudaDeviceSynchronize());               // the table is replaced: no query may be in flight
    if (e->d_links_rows) cudaFree(e->d_links_rows);
    e->d_links_rows = nullptr; e->n_links_rows = 0;
    IGS_CUDA(e, cudaMalloc(&e->d_links_rows, std::max<size_t>(n_rows, 1) * sizeof(igs_links_row)));
    if (n_rows) k_lf_init<<<(n_rows + kThreads - 1) / kThreads, kThreads, 0, st>>>(e->d_links_rows, n_rows);
    IGS_LAUNCHED(e);
    // The links buffer may be stage 2's output (one shard), which lives in the pool: take scratch behind it.
    Scratch sc(e->links_pool);
    if (e->links_pool.size() < kStage3First) e->links_pool.resize(kStage3First);
    sc.next = kStage3First;
    Rec* d_distinct; uint32_t m = 0;
    int rc = distinct_records(e, sc, reinterpret_cast<const Rec*>(d_links), n, &d_distinct, &m, st);
    if (rc != IGS_OK) return rc;
    if (m && n_rows) {
        const uint32_t mb = (m + kThreads - 1) / kThreads;
        for (int round = 0; round < 3; ++round) {
            k_lf_round<<<mb, kThreads, 0, st>>>(d_distinct, m, shard, n_rows, round, e->d_links_rows);
            IGS_LAUNCHED(e);
        }
    }
    if (n_rows) k_lf_close<<<(n_rows + kThreads - 1) / kThreads, kThreads, 0, st>>>(e->d_links_rows, n_rows);
    IGS_LAUNCHED(e);
    IGS_CUDA(e, cudaStreamSynchronize(st));
    e->n_links_rows = n_rows;
    return IGS_OK;     // the stage scratch stays in the pool for the next run of this periodic job
}

int igs_links_finish(igs_engine* e, const igs_link* d_links, uint64_t n64, uint32_t shard) {
    if (!e) return IGS_ERR_INVALID;
    IGS_CUDA(e, cudaSetDevice(e->device));
    std::lock_guard<std::mutex> lk(e->mu);
    return links_finish_impl(e, d_links, n64, shard);
}

// ------------------------------------------------------------------------------------------------
// peer-memory exchange
// ------------------------------------------------------------------------------------------------
static void close_peer_windows(igs_engine* e) {
    for (int s = 0; s < kMaxShards; ++s) {
        if (e->links_peer[s] && e->links_peer_ipc[s]) cudaIpcCloseMemHandle(e->links_peer[s]);
        e->links_peer[s] = nullptr;
        e->links_peer_ipc[s] = false;
    }
}

int igs_links_window_create(igs_engine* e, uint64_t cap_pairs, uint64_t cap_links,
                            uint8_t handle_out[IGS_IPC_HANDLE_BYTES], void** d_window) {
    if (!e) return IGS_ERR_INVALID;
    static_assert(sizeof(cudaIpcMemHandle_t) == IGS_IPC_HANDLE_BYTES, "handle size");
    IGS_REQUIRE(e, cap_pairs < 0xFFFFFFFFull && cap_links < 0xFFFFFFFFull, "window capacities are limited to 2^32-2 records");
    IGS_CUDA(e, cudaSetDevice(e->device));
    std::lock_guard<std::mutex> lk(e->mu);
    IGS_CUDA(e, cudaDeviceSynchronize());
    close_peer_windows(e);
    if (e->links_window) cudaFree(e->links_window);
    e->links_window = nullptr;
    const size_t bytes = kWindowHeader + (size_t)(cap_pairs + cap_links) * sizeof(Rec);
    cudaError_t st = cudaMalloc(&e->links_window, bytes);
    if (st != cudaSuccess) {
        igs_set_error(e, "cudaMalloc(%zu) for the links window failed: %s", bytes, cudaGetErrorString(st));
        return IGS_ERR_NOMEM;
    }
    IGS_CUDA(e, cudaMemset(e->links_window, 0, kWindowHeader));
    e->links_cap_pairs = cap_pairs;
    e->links_cap_links = cap_links;
    if (handle_out) {
        cudaIpcMemHandle_t h;
        IGS_CUDA(e, cudaIpcGetMemHandle(&h, e->links_window));
        memcpy(handle_out, &h, sizeof(h));
    }
    if (d_window) *d_window = e->links_window;
    return IGS_OK;
}

int igs_links_window_open(igs_engine* e, uint32_t peer_shard, const uint8_t handle[IGS_IPC_HANDLE_BYTES]) {
    if (!e) return IGS_ERR_INVALID;
    IGS_REQUIRE(e, peer_shard < kMaxShards && handle, "peer shard out of range / handle is null");
    IGS_CUDA(e, cudaSetDevice(e->device));
    std::lock_guard<std::mutex> lk(e->mu);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    IGS_CUDA(e, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    if (e->links_peer[peer_shard] && e->links_peer_ipc[peer_shard]) cudaIpcCloseMemHandle(e->links_peer[peer_shard]);
    e->links_peer[peer_shard] = p;
    e->links_peer_ipc[peer_shard] = true;
    return IGS_OK;
}

int igs_links_window_attach(igs_engine* e, uint32_t peer_shard, void* d_window) {
    if (!e) return IGS_ERR_INVALID;
    IGS_REQUIRE(e, peer_shard < kMaxShards && d_window, "peer shard out of range / window is null");
    std::lock_guard<std::mutex> lk(e->mu);
    if (e->links_peer[peer_shard] && e->links_peer_ipc[peer_shard]) cudaIpcCloseMemHandle(e->links_peer[peer_shard]);
    e->links_peer[peer_shard] = d_window;
    e->links_peer_ipc[peer_shard] = false;
    return IGS_OK;
}

int igs_links_window_reset(igs_engine* e) {
    if (!e) return IGS_ERR_INVALID;
    if (!e->links_window) { igs_set_error(e, "no links window (igs_links_window_create)"); return IGS_ERR_STATE; }
    IGS_CUDA(e, cudaSetDevice(e->device));
    IGS_CUDA(e, cudaMemsetAsync(e->links_window, 0, kWindowHeader, e->slot[0].stream));
    IGS_CUDA(e, cudaStreamSynchronize(e->slot[0].stream));
    return IGS_OK;
}

// the n_shards windows as a device table (scratch slot 0 of the links pool)
static int upload_windows(igs_engine* e, Scratch& sc, uint32_t n_shards, PeerWin** d_wins, cudaStream_t st) {
    PeerWin h[kMaxShards];
    for (uint32_t s = 0; s < n_shards; ++s) {
        if (!e->links_peer[s]) {
            igs_set_error(e, "the window of shard %u is not attached / opened", s);
            return IGS_ERR_STATE;
        }
        h[s] = window_view(e->links_peer[s], e->links_cap_pairs);
    }
    IGS_CUDA(e, sc.alloc(d_wins, kMaxShards));
    IGS_CUDA(e, cudaMemcpyAsync(*d_wins, h, n_shards * sizeof(PeerWin), cudaMemcpyHostToDevice, st));
    return IGS_OK;
}

int igs_links_pairs_peer(igs_engine* e, uint32_t shard, uint32_t n_shards) {
    if (!e) return IGS_ERR_INVALID;
    IGS_REQUIRE(e, n_shards >= 1 && n_shards <= kMaxShards && shard < n_shards, "shard / n_shards out of range");
    if (!e->has_events || !e->has_event_ids) {
        igs_set_error(e, "no event history with identifiers loaded (igs_load_events_ids)");
        return IGS_ERR_STATE;
    }
    if (!e->links_window) { igs_set_error(e, "no links window (igs_links_window_create)"); return IGS_ERR_STATE; }
    IGS_CUDA(e, cudaSetDevice(e->device));
    std::lock_guard<std::mutex> lk(e->mu);
    cudaStream_t st = e->slot[0].stream;
    Scratch sc(e->links_pool);
    PeerWin* d_wins;
    int rc = upload_windows(e, sc, n_shards, &d_wins, st);
    if (rc != IGS_OK) return rc;
    const uint32_t n = (uint32_t)e->n_raw;
    if (n) {
        k_lp_scatter_peer<<<(n + kThreads * kPeerItems - 1) / (kThreads * kPeerItems), kThreads, 0, st>>>(
            e->raw_device, e->raw_account, n, shard, n_shards, d_wins, (unsigned long long)e->links_cap_pairs);
        IGS_LAUNCHED(e);
    }
    IGS_CUDA(e, cudaStreamSynchronize(st));
    return IGS_OK;
}

// number of records the peers have put into this shard's window (which = 0 pairs, 1 links)
static int window_fill(igs_engine* e, int which, uint64_t cap, uint64_t* n_out) {
    unsigned long long cur[2] = {0, 0};
    IGS_CUDA(e, cudaMemcpy(cur, e->links_window, sizeof(cur), cudaMemcpyDeviceToHost));
    if (cur[which] > cap) {
        igs_set_error(e, "links window too small: %llu %s arrived, capacity %llu (igs_links_window_create)",
                      cur[which], which ? "links" : "pairs", (unsigned long long)cap);
        return IGS_ERR_NOMEM;
    }
    *n_out = cur[which];
    return IGS_OK;
}

int igs_links_group_peer(igs_engine* e, uint32_t shard, uint32_t n_shards) {
    if (!e) return IGS_ERR_INVALID;
    IGS_REQUIRE(e, n_shards >= 1 && n_shards <= kMaxShards && shard < n_shards, "shard / n_shards out of range");
    if (!e->links_window) { igs_set_error(e, "no links window (igs_links_window_create)"); return IGS_ERR_STATE; }
    IGS_CUDA(e, cudaSetDevice(e->device));
    std::lock_guard<std::mutex> lk(e->mu);
    uint64_t n = 0;
    int rc = window_fill(e, 0, e->links_cap_pairs, &n);
    if (rc != IGS_OK) return rc;
    Scratch sc(e->links_pool);
    PeerWin* d_wins;
    rc = upload_windows(e, sc, n_shards, &d_wins, e->slot[0].stream);
    if (rc != IGS_OK) return rc;
    const PeerWin mine = window_view(e->links_window, e->links_cap_pairs);
    return links_group_impl(e, reinterpret_cast<const igs_link_pair*>(mine.pairs), n, n_shards, nullptr, nullptr, d_wins);
}

int igs_links_finish_peer(igs_engine* e, uint32_t shard) {
    if (!e) return IGS_ERR_INVALID;
    if (!e->links_window) { igs_set_error(e, "no links window (igs_links_window_create)"); return IGS_ERR_STATE; }
    IGS_CUDA(e, cudaSetDevice(e->device));
    std::lock_guard<std::mutex> lk(e->mu);
    uint64_t n = 0;
    int rc = window_fill(e, 1, e->links_cap_links, &n);
    if (rc != IGS_OK) return rc;
    const PeerWin mine = window_view(e->links_window, e->links_cap_pairs);
    return links_finish_impl(e, reinterpret_cast<const igs_link*>(mine.links), n, shard);
}

int igs_get_links(igs_engine* e, const uint32_t* accounts, uint64_t n, igs_links_row* out) {
    if (!e) return IGS_ERR_INVALID;
    IGS_REQUIRE(e, n == 0 || (accounts && out), "accounts / out is null");
    if (n == 0) return IGS_OK;
    IGS_CUDA(e, cudaSetDevice(e->device));
    igs_slot& sl = e->slot[0];
    int rc = igs_reserve(e, &sl.in0, n * 4);
    if (rc != IGS_OK) return rc;
    rc = igs_reserve(e, &sl.out0, n * sizeof(igs_links_row));
    if (rc != IGS_OK) return rc;
    IGS_CUDA(e, cudaMemcpyAsync(sl.in0.p, accounts, n * 4, cudaMemcpyHostToDevice, sl.stream));
    k_links_get<<<(unsigned)((n + kThreads - 1) / kThreads), kThreads, 0, sl.stream>>>(
        e->d_links_rows, e->n_links_rows, (const uint32_t*)sl.in0.p, n, (igs_links_row*)sl.out0.p);
    IGS_LAUNCHED(e);
    IGS_CUDA(e, cudaMemcpyAsync(out, sl.out0.p, n * sizeof(igs_links_row), cudaMemcpyDeviceToHost, sl.stream));
    IGS_CUDA(e, cudaStreamSynchronize(sl.stream));
    return IGS_OK;
}

int igs_check_bonus_abuse(igs_engine* e, const uint32_t* accounts, uint64_t n, igs_abuse* out) {
    if (!e) return IGS_ERR_INVALID;
    IGS_REQUIRE(e, n == 0 || (accounts && out), "accounts / out is null");
    if (n == 0) return IGS_OK;
    IGS_CUDA(e, cudaSetDevice(e->device));
    igs_slot& sl = e->slot[0];
    int rc = igs_reserve(e, &sl.in0, n * 4);
    if (rc != IGS_OK) return rc;
    rc = igs_reserve(e, &sl.out0, n * sizeof(igs_abuse));
    if (rc != IGS_OK) return rc;
    IGS_CUDA(e, cudaMemcpyAsync(sl.in0.p, accounts, n * 4, cudaMemcpyHostToDevice, sl.stream));
    k_bonus_abuse<<<(unsigned)((n + kThreads - 1) / kThreads), kThreads, 0, sl.stream>>>(
        e->d_links_rows, e->n_links_rows, e->d_players, e->d_players ? e->n_players : 0, (const uint32_t*)sl.in0.p, n,
        (igs_abuse*)sl.out0.p);
    IGS_LAUNCHED(e);
    IGS_CUDA(e, cudaMemcpyAsync(out, sl.out0.p, n * sizeof(igs_abuse), cudaMemcpyDeviceToHost, sl.stream));
    IGS_CUDA(e, cudaStreamSynchronize(sl.stream));
    return IGS_OK;
}

int igs_abuse_signals(uint32_t signals, const char* out[3]) {
    static const char* names[3] = {"BONUS_ONLY_PLAYER", "MULTI_ACCOUNT", "ACCOUNT_RING"};
    int k = 0;
    for (int b = 0; b < 3; ++b)
        if (signals & (1u << b)) out[k++] = names[b];
    return k;
}

}  // extern "C"
