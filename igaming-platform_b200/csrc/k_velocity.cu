// k_velocity.cu — event-history index: radix sort + segmented scans (K3) + the
// per-query velocity kernel.  sm_100a; all kernels are HBM-streaming integer work.
//
// Build (igs_load_events), replaces the Redis state of UpdateRealTimeFeatures
// (features/redis_store.go:119-168):
//   1. key = (account << ts_bits) | (ts - ts_min), value = arrival index; LSD radix sort, stable,
//      <= 8-bit digits sized so that the key's significant bits take the fewest passes
//      (ZSET ordering by score, :127-130).
//   2. gather amount / type by the sorted arrival index; CSR `seg` of account boundaries.
//   3. distinct flag per event: first occurrence of (account, ts, amount) — the ZSET member
//      "ts:amount" dedupe (:129).
//   4. one fused scan over the sorted events producing, per position, the exclusive prefix
//      {distinct count, per-type count, per-type sum} and, per event, the inclusive segmented
//      {chain sum with head = new account || gap > 3600, session start with head = new account ||
//      gap > 1800}.  A 256-ary hierarchy of warp-shuffle passes (reduce up, scan the top, apply down).
//   5. (histories that carry identifiers, igs_load_events_ids) per identifier column — device, IP — the
//      PFADD + EXPIRE 24h key of :141-152 as of every event: compact the events that carry an identifier,
//      cut them into chains (gap > 86400 or new account), find each event's previous occurrence of the
//      same identifier in its chain through a stable radix sort on a 32-bit hash of the identifier
//      (exact: equal hashes are resolved by comparing the 64-bit ids), and count first occurrences per
//      chain with a prefix sum.
// Query (igs_get_velocity*, and inlined into the fused score kernel): velocity.cuh.
#include <string.h>

#include <algorithm>
#include <vector>

#include "idset.cuh"
#include "igs_internal.cuh"
#include "onesweep.cuh"
#include "sortscan.cuh"
#include "velocity.cuh"

namespace {

// ------------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------------

__global__ void k_minmax(const uint32_t* __restrict__ account, const uint32_t* __restrict__ ts,
                         uint32_t n, uint32_t* __restrict__ out /* ts_min, ts_max, acct_max, unsorted */) {
    uint32_t tmin = 0xFFFFFFFFu, tmax = 0, amax = 0, unsorted = 0;
    // four events per step from the first 16-byte boundary on (both columns start at the same event offset of their
    // allocation, so they share it); the up to 3 + 3 events before and after go to block 0's first threads
    const uint32_t head = min(n, (uint32_t)((16u - ((uintptr_t)ts & 15u)) & 15u) >> 2);
    const uint4* ts4 = reinterpret_cast<const uint4*>(ts + head);
    const uint4* ac4 = reinterpret_cast<const uint4*>(account + head);
    const uint32_t n4 = (n - head) >> 2;
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n4; j += gridDim.x * blockDim.x) {
        const uint4 t = __ldg(ts4 + j), a = __ldg(ac4 + j);
        const uint32_t i = head + 4 * j;
        const uint32_t prev = i ? __ldg(ts + i - 1) : 0u;
        tmin = min(min(tmin, min(t.x, t.y)), min(t.z, t.w));
        tmax = max(max(tmax, max(t.x, t.y)), max(t.z, t.w));
        amax = max(max(amax, max(a.x, a.y)), max(a.z, a.w));
        if (prev > t.x || t.x > t.y || t.y > t.z || t.z > t.w) unsorted = 1;    // arrival order is not time order
    }
    const uint32_t tail0 = head + (n4 << 2), n_edge = head + (n - tail0);
    if (blockIdx.x == 0 && threadIdx.x < n_edge) {
        const uint32_t i = threadIdx.x < head ? threadIdx.x : tail0 + (threadIdx.x - head);
        const uint32_t t = __ldg(ts + i), a = __ldg(account + i);
        tmin = min(tmin, t); tmax = max(tmax, t); amax = max(amax, a);
        if (i > 0 && __ldg(ts + i - 1) > t) unsorted = 1;
    }
    if (__any_sync(0xffffffffu, unsorted) && (threadIdx.x & 31) == 0) atomicOr(out + 3, 1u);
    for (int o = 16; o; o >>= 1) {
        tmin = min(tmin, __shfl_xor_sync(0xffffffffu, tmin, o));
        tmax = max(tmax, __shfl_xor_sync(0xffffffffu, tmax, o));
        amax = max(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(out + 0, tmin);
        atomicMax(out + 1, tmax);
        atomicMax(out + 2, amax);
    }
}

__global__ void k_make_keys(const uint32_t* __restrict__ account, const uint32_t* __restrict__ ts,
                            uint32_t n, uint32_t ts_min, int ts_bits, uint64_t* __restrict__ keys,
                            uint32_t* __restrict__ vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    keys[i] = ts_bits ? ((uint64_t)__ldg(account + i) << ts_bits) | (uint64_t)(__ldg(ts + i) - ts_min)
                      : (uint64_t)__ldg(account + i);       // time-ordered arrival: a stable sort by account is enough
    vals[i] = i;
}

// ---- packed keys: everything the sort has to carry in one 64-bit word --------------------------------------
//   [ account | ts - ts_min | type (2 bits) | arrival index inside the range ]      (high ... low)
// sorted on the account (and ts) bits only; the low fields ride along, so that after the sort the account, the
// timestamp, the type and the arrival index come out of the key and only the amount has to be gathered.
// ts_bits == 0: the events are already in time order per account (a log, micro-batches of a live feed): a stable
// sort by account keeps that order and the timestamp is gathered like the amount.
struct KeyLayout {
    int seq_bits, ts_bits, acct_shift;     // acct_shift = seq_bits + 2 + ts_bits
    uint32_t ts_min;
};
// Builds the keys and, in the same pass over the raw columns, the global digit histograms of every sort pass.
// bad[0] = largest account index seen that is >= n_accounts (0 = none).
__global__ void __launch_bounds__(kOsThreads)
k_make_keys_hist(const uint32_t* __restrict__ account, const uint32_t* __restrict__ ts,
                 const uint8_t* __restrict__ type, uint32_t n, uint32_t n_accounts, KeyLayout lay, OsPlan plan,
                 OsScratch sc, uint64_t* __restrict__ keys, uint32_t* __restrict__ bad) {
    extern __shared__ uint32_t mk_sh_hist[];
    __shared__ OsPlan sh_plan;                     // indexed per pass: shared memory, not a local copy of the parameter
    if (threadIdx.x == 0) sh_plan = plan;
    os_hist_clear(mk_sh_hist, plan.passes);
    __syncthreads();
    uint32_t worst = 0;
    constexpr int kU = 4;                          // four events per thread and round: their loads are in flight together
    const uint32_t stride = gridDim.x * kOsThreads;
    for (uint32_t i0 = blockIdx.x * kOsThreads + threadIdx.x; i0 < n; i0 += stride * kU) {
        uint32_t a[kU], t[kU], ty[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const uint32_t i = i0 + u * stride;
            const bool ok = i < n;
            a[u] = ok ? __ldg(account + i) : 0u;
            t[u] = ok && lay.ts_bits ? __ldg(ts + i) : lay.ts_min;
            ty[u] = ok ? min((uint32_t)__ldg(type + i), 3u) : 0u;
        }
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const uint32_t i = i0 + u * stride;
            if (i >= n) break;
            if (a[u] >= n_accounts) worst = max(worst, a[u]);
            const uint64_t k = ((uint64_t)a[u] << lay.acct_shift) | ((uint64_t)(t[u] - lay.ts_min) << (lay.seq_bits + 2)) |
                               ((uint64_t)ty[u] << lay.seq_bits) | (uint64_t)i;
            keys[i] = k;
            os_hist_accumulate(mk_sh_hist, sh_plan, k);
        }
    }
    if (worst) atomicMax(bad, worst);
    __syncthreads();
    os_hist_flush(mk_sh_hist, sh_plan, sc);
}

__global__ void k_unpack_gather_packed(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ ts_in,
                                       const int64_t* __restrict__ amount_in, uint32_t n, KeyLayout lay, uint32_t seq_base,
                                       uint32_t* __restrict__ acct, uint32_t* __restrict__ ts, uint32_t* __restrict__ seq_out,
                                       int64_t* __restrict__ amount, uint8_t* __restrict__ type) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t k = __ldg(keys + i);
    const uint32_t local = (uint32_t)(k & ((1ull << lay.seq_bits) - 1ull));
    acct[i] = (uint32_t)(k >> lay.acct_shift);
    type[i] = (uint8_t)((k >> lay.seq_bits) & 3u);
    ts[i] = lay.ts_bits ? lay.ts_min + (uint32_t)((k >> (lay.seq_bits + 2)) & ((1ull << lay.ts_bits) - 1ull))
                        : __ldg(ts_in + local);
    seq_out[i] = local + seq_base;
    amount[i] = __ldg(amount_in + local);
}

// ------------------------------------------------------------------------------------------------
// post-sort columns
// ------------------------------------------------------------------------------------------------
// `local` = position of the sorted event inside the raw range being indexed (the *_in columns start at the range);
// seq_out receives its global arrival index (local + seq_base): base and delta share one numbering.
__global__ void k_unpack_gather(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ local,
                                const uint32_t* __restrict__ ts_in,
                                const int64_t* __restrict__ amount_in, const uint8_t* __restrict__ type_in,
                                uint32_t n, uint32_t ts_min, int ts_bits, uint32_t seq_base,
                                uint32_t* __restrict__ acct, uint32_t* __restrict__ ts, uint32_t* __restrict__ seq_out,
                                int64_t* __restrict__ amount, uint8_t* __restrict__ type) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t k = __ldg(keys + i);
    acct[i] = (uint32_t)(k >> ts_bits);
    const uint32_t s = __ldg(local + i);
    seq_out[i] = s + seq_base;
    ts[i] = ts_bits ? ts_min + (uint32_t)(k & ((1ull << ts_bits) - 1ull)) : __ldg(ts_in + s);
    amount[i] = __ldg(amount_in + s);
    type[i] = __ldg(type_in + s);
}

// CSR segments + (account, ts) group heads + LATE events, one pass over the sorted columns.
// Late: sorted neighbours of one account whose arrival order contradicts their time order — either the earlier-
// timestamped one arrived later (sort by (account, ts): arrival indices invert), or, under a sort by account alone,
// the later arrival carries the smaller timestamp.  Redis would have applied a late event on arrival, with its old
// timestamp as "now" (features/redis_store.go:121,132,155); the index files events where their timestamps say, and
// under an account-only sort a late pair means the order is not the timestamp order at all (the caller rejects or
// re-sorts).  late[0] = late pairs in the range, late[1] = those whose late event has arrival index >= n_before
// (the batch being appended).  When a delta is built over a base (base.n != 0), the first delta event of an
// account is also late if it is older than the account's last base event.
__global__ void __launch_bounds__(kThreads)
k_segments_heads(const uint32_t* __restrict__ acct, const uint32_t* __restrict__ ts,
                 const uint32_t* __restrict__ seq, uint32_t n, uint32_t n_accounts, uint32_t n_before,
                 igs_index_view base, uint32_t* __restrict__ seg, uint8_t* __restrict__ head,
                 uint32_t* __restrict__ late) {
    __shared__ uint32_t sh_late[2];
    if (threadIdx.x < 2) sh_late[threadIdx.x] = 0;
    __syncthreads();
    uint32_t n_late = 0, n_new = 0;                // a history in random time order is full of late pairs: count locally
    for (uint32_t i = blockIdx.x * kThreads + threadIdx.x; i < n; i += gridDim.x * kThreads) {
        // (an out-of-range account is reported by the key pass; clamped here so that seg[] is never overrun)
        const uint32_t a = min(__ldg(acct + i), n_accounts), t = __ldg(ts + i);
        bool first_of_account = i == 0, group_head = i == 0;
        bool is_late = false, is_new = false;
        if (i == 0) {
            for (uint32_t x = 0; x <= a; ++x) seg[x] = 0;
        } else {
            const uint32_t p = min(__ldg(acct + i - 1), n_accounts), tp = __ldg(ts + i - 1);
            for (uint32_t x = p + 1; x <= a; ++x) seg[x] = i;
            first_of_account = p != a;
            group_head = first_of_account || tp != t;
            if (!first_of_account) {
                const uint32_t sp = __ldg(seq + i - 1), si = __ldg(seq + i);
                if (sp > si) { is_late = true; is_new = sp >= n_before; }          // i-1 arrived later, is not newer
                else if (tp > t) { is_late = true; is_new = si >= n_before; }       // i arrived later, is older
            }
        }
        head[i] = group_head;
        if (first_of_account && base.n != 0 && a < base.n_accounts) {
            const uint32_t sb = __ldg(base.seg + a), eb = __ldg(base.seg + a + 1);
            if (eb > sb && t < __ldg(base.ts + eb - 1)) {
                is_late = true;
                is_new = __ldg(seq + i) >= n_before;
            }
        }
        if (i == n - 1)
            for (uint32_t x = a + 1; x <= n_accounts; ++x) seg[x] = n;
        n_late += is_late;
        n_new += is_new;
    }
    n_late = __reduce_add_sync(0xffffffffu, n_late);
    n_new = __reduce_add_sync(0xffffffffu, n_new);
    if ((threadIdx.x & 31) == 0 && n_late) { atomicAdd(&sh_late[0], n_late); atomicAdd(&sh_late[1], n_new); }
    __syncthreads();
    if (threadIdx.x == 0 && sh_late[0]) { atomicAdd(late + 0, sh_late[0]); atomicAdd(late + 1, sh_late[1]); }
}

// distinct flag: no earlier event of the same (account, ts) group has the same amount (the ZSET member
// "ts:amount" dedupe).  Walks back through the group four events at a time (the loads of a step do not
// depend on each other), at most kDistinctWalk events: groups of up to kDistinctWalk + 1 events — all of
// them in ordinary traffic — are settled here.  The members of LONGER groups (a bot, a replayed queue: a
// same-second burst of any size) are collected in a work list instead and settled by the hash pass below, so
// that such a burst costs linear, not quadratic, work — and only its own members pay for it.
constexpr uint32_t kDistinctWalk = 32;

__global__ void k_distinct(const uint8_t* __restrict__ head, const int64_t* __restrict__ amount, uint32_t n,
                           uint8_t* __restrict__ dflag, uint32_t* __restrict__ work, uint32_t* __restrict__ n_work) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint8_t d = 1;
    bool big = false;
    if (i < n) {
        uint32_t start = i;                       // group head, once known
        bool found_head = __ldg(head + i) != 0;
        if (!found_head) {
            const int64_t m = __ldg(amount + i);
            uint32_t j = i;                       // candidates are j-1, j-2, ... down to the group head
            bool dup = false;
            for (uint32_t step = 0; step < kDistinctWalk / 4 && !found_head; ++step) {
                int64_t a[4];
                uint8_t h[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint32_t jj = j >= (uint32_t)(k + 1) ? j - (k + 1) : 0;
                    a[k] = __ldg(amount + jj);
                    h[k] = j >= (uint32_t)(k + 1) ? __ldg(head + jj) : 1;
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (found_head) break;
                    if (a[k] == m) dup = true;
                    if (h[k]) { found_head = true; start = j - (k + 1); }   // reached the first event of the group
                }
                j = j >= 4 ? j - 4 : 0;
            }
            d = dup ? 0 : 1;                      // (a duplicate seen before the head was reached is a duplicate either way)
        }
        if (!found_head) {
            big = true;                           // deeper than the walk: the group has more than kDistinctWalk + 1 members
        } else {
            // a member near the head of a long group must be in the hash table too: does the group reach past
            // start + kDistinctWalk + 1 ?
            big = true;
            for (uint32_t j = i + 1; j <= start + kDistinctWalk + 1; ++j)
                if (j >= n || __ldg(head + j)) { big = false; break; }
        }
        dflag[i] = d;
    }
    const unsigned m = __ballot_sync(0xffffffffu, big);
    if (m) {
        const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
        uint32_t base = 0;
        if (lane == leader) base = atomicAdd(n_work, (uint32_t)__popc(m));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (big) work[base + __popc(m & ((1u << lane) - 1u))] = i;
    }
}

// ---- hash pass for the long groups (rare) ------------------------------------------------------------
// One open-addressing table for all of them.  A slot is claimed by the first member that gets there and holds that
// member's sorted position (its "representative"); later members compare their (account, ts, amount) with the
// representative's — exact, whatever the hash does — and the smallest position per (account, ts, amount) wins.
__device__ __forceinline__ uint32_t dg_hash(uint32_t a, uint32_t t, int64_t m, uint32_t mask) {
    return (uint32_t)(igs_idset_mix(((uint64_t)a << 32 | t) ^ igs_idset_mix((uint64_t)m)) & mask);
}
__global__ void k_dg_insert(const uint32_t* __restrict__ work, uint32_t n_work, const uint32_t* __restrict__ acct,
                            const uint32_t* __restrict__ ts, const int64_t* __restrict__ amount, uint32_t mask,
                            uint32_t* __restrict__ rep, uint32_t* __restrict__ minpos) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_work) return;
    const uint32_t i = __ldg(work + k);
    const uint32_t a = __ldg(acct + i), t = __ldg(ts + i);
    const int64_t m = __ldg(amount + i);
    uint32_t h = dg_hash(a, t, m, mask);
    for (;;) {
        uint32_t r = atomicCAS(rep + h, 0xFFFFFFFFu, i);
        if (r == 0xFFFFFFFFu) r = i;
        if (r == i || (__ldg(acct + r) == a && __ldg(ts + r) == t && __ldg(amount + r) == m)) {
            atomicMin(minpos + h, i);
            return;
        }
        h = (h + 1) & mask;
    }
}
__global__ void k_dg_resolve(const uint32_t* __restrict__ work, uint32_t n_work, const uint32_t* __restrict__ acct,
                             const uint32_t* __restrict__ ts, const int64_t* __restrict__ amount, uint32_t mask,
                             const uint32_t* __restrict__ rep, const uint32_t* __restrict__ minpos,
                             uint8_t* __restrict__ dflag) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_work) return;
    const uint32_t i = __ldg(work + k);
    const uint32_t a = __ldg(acct + i), t = __ldg(ts + i);
    const int64_t m = __ldg(amount + i);
    uint32_t h = dg_hash(a, t, m, mask);
    for (;;) {
        const uint32_t r = rep[h];                // present: k_dg_insert put this member's key there
        if (r == i || (__ldg(acct + r) == a && __ldg(ts + r) == t && __ldg(amount + r) == m)) {
            dflag[i] = minpos[h] == i;
            return;
        }
        h = (h + 1) & mask;
    }
}

// ---- delta over a base: the ZSET member dedupe across the boundary --------------------------------------
// A delta event in the first (account, ts) group of its account whose ts equals the ts of the account's last
// base event shares that second with base events: it is not a first occurrence if a base event of that second
// carries the same amount.  (Later delta groups have a larger ts than every base event.)  Linear walk over the
// base's trailing group — a same-second burst that straddles an append.
__global__ void k_delta_boundary_dedupe(const uint32_t* __restrict__ acct, const uint32_t* __restrict__ ts,
                                        const int64_t* __restrict__ amount, const uint32_t* __restrict__ seg,
                                        uint32_t n, igs_index_view base, const int64_t* __restrict__ raw_amount,
                                        uint8_t* __restrict__ dflag) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !dflag[i]) return;
    const uint32_t a = __ldg(acct + i);
    if (a >= base.n_accounts) return;
    const uint32_t sb = __ldg(base.seg + a), eb = __ldg(base.seg + a + 1);
    if (eb == sb) return;
    const uint32_t t = __ldg(ts + i);
    if (__ldg(base.ts + eb - 1) != t || __ldg(ts + __ldg(seg + a)) != t) return;
    const int64_t m = __ldg(amount + i);
    for (uint32_t k = eb; k-- > sb && __ldg(base.ts + k) == t;)
        if (__ldg(raw_amount + __ldg(base.seq + k)) == m) { dflag[i] = 0; return; }
}

// ------------------------------------------------------------------------------------------------
// fused scan of the velocity columns
// ------------------------------------------------------------------------------------------------
struct VScan {
    int64_t  s0, s1, chain;
    uint32_t d, c0, c1, st;
    uint32_t cf, sf;   // segment-head flags seen (chain / session)
};

__device__ __forceinline__ VScan vs_identity() {
    VScan r; r.s0 = r.s1 = r.chain = 0; r.d = r.c0 = r.c1 = r.st = 0; r.cf = r.sf = 0; return r;
}
__device__ __forceinline__ VScan vs_combine(const VScan& a, const VScan& b) {   // a earlier
    VScan r;
    r.d = a.d + b.d; r.c0 = a.c0 + b.c0; r.c1 = a.c1 + b.c1;
    r.s0 = a.s0 + b.s0; r.s1 = a.s1 + b.s1;
    r.chain = b.cf ? b.chain : a.chain + b.chain; r.cf = a.cf | b.cf;
    r.st = b.sf ? b.st : a.st; r.sf = a.sf | b.sf;
    return r;
}
struct VCols {
    const uint32_t* acct; const uint32_t* ts; const int64_t* amount; const uint8_t* type;
    const uint8_t* dflag; uint32_t n;
    igs_index_view base;       // base.n != 0: a delta is being built and the chains continue the base's
};
__device__ __forceinline__ VScan vs_element(const VCols& c, uint32_t i) {
    VScan r = vs_identity();
    const uint32_t t = __ldg(c.ts + i);
    const int64_t m = __ldg(c.amount + i);
    const uint8_t ty = __ldg(c.type + i);
    bool new_acct = i == 0;
    long long gap = 0;
    if (i > 0) {
        new_acct = __ldg(c.acct + i) != __ldg(c.acct + i - 1);
        gap = (long long)t - (long long)__ldg(c.ts + i - 1);
    }
    r.d = __ldg(c.dflag + i);
    if (ty == IGS_TX_BET) { r.c0 = 1; r.s0 = m; }
    if (ty == IGS_TX_DEPOSIT) { r.c1 = 1; r.s1 = m; }
    r.chain = m;
    r.cf = (new_acct || gap > 3600) ? 1u : 0u;        // tx_sum:1h key expired (redis_store.go:138)
    r.sf = (new_acct || gap > 1800) ? 1u : 0u;        // session key expired (:159-160)
    r.st = r.sf ? t : 0u;
    if (new_acct && c.base.n != 0) {
        // first delta event of the account: the keys it finds are the ones the account's last base event left
        const uint32_t a = __ldg(c.acct + i);
        if (a < c.base.n_accounts) {
            const uint32_t sb = __ldg(c.base.seg + a), eb = __ldg(c.base.seg + a + 1);
            if (eb > sb) {
                const long long gb = (long long)t - (long long)__ldg(c.base.ts + eb - 1);
                const igs_tail_rec tb = c.base.tail[eb - 1];
                if (gb <= 3600) r.chain = m + tb.chain_sum;      // still a segment head for the scan: the value is complete
                if (gb <= 1800) r.st = tb.sess_start;
            }
        }
    }
    return r;
}

// The scan is a 256-ary hierarchy of warp-level passes (no shared memory, no block barriers): a warp owns
// 256 consecutive items, reads them 32 at a time (coalesced) and reduces / scans each round with shuffles,
// carrying the running value in registers.
//   level 0: elements        -> one aggregate per 256 events        (k_vs_reduce<ElemSrc>)
//   level 1: those           -> one aggregate per 65 536 events      (k_vs_reduce<AggSrc>)
//   level 2: a single warp scans the few level-1 aggregates          (k_vs_scan_small)
//   down   : exclusive scan of the level-0 aggregates with carry-in  (k_vs_apply_aggs),
//            then of the elements, writing the prefix / tail records (k_vs_apply_elems)
constexpr int kVTile = 256;

__device__ __forceinline__ VScan vs_shfl_up(const VScan& v, int o) {
    VScan r;
    r.s0 = __shfl_up_sync(0xffffffffu, v.s0, o); r.s1 = __shfl_up_sync(0xffffffffu, v.s1, o);
    r.chain = __shfl_up_sync(0xffffffffu, v.chain, o);
    r.d = __shfl_up_sync(0xffffffffu, v.d, o); r.c0 = __shfl_up_sync(0xffffffffu, v.c0, o);
    r.c1 = __shfl_up_sync(0xffffffffu, v.c1, o); r.st = __shfl_up_sync(0xffffffffu, v.st, o);
    const uint32_t f = __shfl_up_sync(0xffffffffu, v.cf | (v.sf << 1), o);
    r.cf = f & 1u; r.sf = f >> 1;
    return r;
}
__device__ __forceinline__ VScan vs_shfl(const VScan& v, int src) {
    VScan r;
    r.s0 = __shfl_sync(0xffffffffu, v.s0, src); r.s1 = __shfl_sync(0xffffffffu, v.s1, src);
    r.chain = __shfl_sync(0xffffffffu, v.chain, src);
    r.d = __shfl_sync(0xffffffffu, v.d, src); r.c0 = __shfl_sync(0xffffffffu, v.c0, src);
    r.c1 = __shfl_sync(0xffffffffu, v.c1, src); r.st = __shfl_sync(0xffffffffu, v.st, src);
    const uint32_t f = __shfl_sync(0xffffffffu, v.cf | (v.sf << 1), src);
    r.cf = f & 1u; r.sf = f >> 1;
    return r;
}
// inclusive scan across the 32 lanes (lane order = item order)
__device__ __forceinline__ VScan vs_warp_scan(VScan v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const VScan y = vs_shfl_up(v, o);
        if (lane >= o) v = vs_combine(y, v);
    }
    return v;
}

struct AggSrc {
    const VScan* a; uint32_t n;
    __device__ __forceinline__ VScan operator()(uint32_t i) const { return a[i]; }
    __device__ __forceinline__ uint32_t size() const { return n; }
};

// Level 0 works lane-serially: a lane owns 8 consecutive events (vector loads: 32 B of accounts, 32 B of
// timestamps, 64 B of amounts, 8 B of types, 8 B of flags), combines them in order in registers, and the warp
// scans the 32 lane aggregates once — one round of shuffles per 256 events instead of eight.
constexpr int kVLane = kVTile / 32;      // 8
static_assert(kVLane == 8, "vector loads below assume 8 events per lane");

struct VLaneItems {
    uint32_t a[kVLane], t[kVLane];
    int64_t  m[kVLane];
    uint8_t  ty[kVLane], d[kVLane];
    uint32_t prev_a, prev_t;             // event first - 1 (undefined when first == 0)
};
__device__ __forceinline__ void vs_load_lane(const VCols& c, uint32_t first, VLaneItems& it) {
    if (first + kVLane <= c.n) {
        const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(c.acct + first)), a1 = __ldg(reinterpret_cast<const uint4*>(c.acct + first) + 1);
        const uint4 t0 = __ldg(reinterpret_cast<const uint4*>(c.ts + first)), t1 = __ldg(reinterpret_cast<const uint4*>(c.ts + first) + 1);
        it.a[0] = a0.x; it.a[1] = a0.y; it.a[2] = a0.z; it.a[3] = a0.w; it.a[4] = a1.x; it.a[5] = a1.y; it.a[6] = a1.z; it.a[7] = a1.w;
        it.t[0] = t0.x; it.t[1] = t0.y; it.t[2] = t0.z; it.t[3] = t0.w; it.t[4] = t1.x; it.t[5] = t1.y; it.t[6] = t1.z; it.t[7] = t1.w;
        const longlong2* mp = reinterpret_cast<const longlong2*>(c.amount + first);
#pragma unroll
        for (int k = 0; k < 4; ++k) { const longlong2 v = __ldg(mp + k); it.m[2 * k] = v.x; it.m[2 * k + 1] = v.y; }
        const uint2 ty = __ldg(reinterpret_cast<const uint2*>(c.type + first));
        const uint2 df = __ldg(reinterpret_cast<const uint2*>(c.dflag + first));
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            it.ty[k] = (uint8_t)(ty.x >> (8 * k)); it.ty[4 + k] = (uint8_t)(ty.y >> (8 * k));
            it.d[k] = (uint8_t)(df.x >> (8 * k)); it.d[4 + k] = (uint8_t)(df.y >> (8 * k));
        }
    } else {
#pragma unroll
        for (int k = 0; k < kVLane; ++k) {
            const uint32_t i = first + k;
            const bool ok = i < c.n;
            it.a[k] = ok ? __ldg(c.acct + i) : 0; it.t[k] = ok ? __ldg(c.ts + i) : 0; it.m[k] = ok ? __ldg(c.amount + i) : 0;
            it.ty[k] = ok ? __ldg(c.type + i) : 0; it.d[k] = ok ? __ldg(c.dflag + i) : 0;
        }
    }
    it.prev_a = first > 0 && first <= c.n ? __ldg(c.acct + first - 1) : 0;
    it.prev_t = first > 0 && first <= c.n ? __ldg(c.ts + first - 1) : 0;
}
// element k of the lane (same arithmetic as vs_element, from registers)
__device__ __forceinline__ VScan vs_lane_element(const VCols& c, const VLaneItems& it, uint32_t first, int k) {
    VScan r = vs_identity();
    const uint32_t i = first + k;
    const uint32_t t = it.t[k];
    const int64_t m = it.m[k];
    const bool new_acct = i == 0 || it.a[k] != (k ? it.a[k - 1] : it.prev_a);
    const long long gap = i == 0 ? 0 : (long long)t - (long long)(k ? it.t[k - 1] : it.prev_t);
    r.d = it.d[k];
    if (it.ty[k] == IGS_TX_BET) { r.c0 = 1; r.s0 = m; }
    if (it.ty[k] == IGS_TX_DEPOSIT) { r.c1 = 1; r.s1 = m; }
    r.chain = m;
    r.cf = (new_acct || gap > 3600) ? 1u : 0u;        // tx_sum:1h key expired (redis_store.go:138)
    r.sf = (new_acct || gap > 1800) ? 1u : 0u;        // session key expired (:159-160)
    r.st = r.sf ? t : 0u;
    if (new_acct && c.base.n != 0) {
        // first delta event of the account: the keys it finds are the ones the account's last base event left
        const uint32_t a = it.a[k];
        if (a < c.base.n_accounts) {
            const uint32_t sb = __ldg(c.base.seg + a), eb = __ldg(c.base.seg + a + 1);
            if (eb > sb) {
                const long long gb = (long long)t - (long long)__ldg(c.base.ts + eb - 1);
                const igs_tail_rec tb = c.base.tail[eb - 1];
                if (gb <= 3600) r.chain = m + tb.chain_sum;      // still a segment head for the scan: the value is complete
                if (gb <= 1800) r.st = tb.sess_start;
            }
        }
    }
    return r;
}

// one aggregate per 256 events
__global__ void __launch_bounds__(kThreads)
k_vs_reduce_elems(VCols c, VScan* __restrict__ agg) {
    const uint32_t gw = (blockIdx.x * kThreads + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t base = gw * kVTile;
    if (base >= c.n) return;
    const uint32_t first = base + lane * kVLane;
    VLaneItems it;
    vs_load_lane(c, first, it);
    VScan v = vs_identity();
#pragma unroll
    for (int k = 0; k < kVLane; ++k)
        if (first + k < c.n) v = vs_combine(v, vs_lane_element(c, it, first, k));
    const VScan inc = vs_warp_scan(v, lane);
    if (lane == 31) agg[gw] = inc;
}

template <typename Src>
__global__ void __launch_bounds__(kThreads)
k_vs_reduce(Src src, VScan* __restrict__ agg) {
    const uint32_t gw = (blockIdx.x * kThreads + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t n = src.size(), base = gw * kVTile;
    if (base >= n) return;
    VScan carry = vs_identity();
#pragma unroll 1
    for (int r = 0; r < kVTile / 32; ++r) {
        const uint32_t i = base + r * 32 + lane;
        const VScan inc = vs_warp_scan(i < n ? src(i) : vs_identity(), lane);
        carry = vs_combine(carry, vs_shfl(inc, 31));
    }
    if (lane == 0) agg[gw] = carry;
}

// exclusive scan of a[0..n) in place by one warp (n is a few hundred at most)
__global__ void __launch_bounds__(32)
k_vs_scan_small(VScan* __restrict__ a, uint32_t n) {
    const int lane = threadIdx.x;
    VScan carry = vs_identity();
    for (uint32_t base = 0; base < n; base += 32) {
        const uint32_t i = base + lane;
        const VScan v = i < n ? a[i] : vs_identity();
        const VScan inc = vs_warp_scan(v, lane);
        VScan ex = vs_shfl_up(inc, 1);
        if (lane == 0) ex = vs_identity();
        if (i < n) a[i] = vs_combine(carry, ex);
        carry = vs_combine(carry, vs_shfl(inc, 31));
    }
}

// exclusive scan of a[0..n) in place, 256 items per warp, carry-in per warp tile from tile_excl
__global__ void __launch_bounds__(kThreads)
k_vs_apply_aggs(VScan* __restrict__ a, uint32_t n, const VScan* __restrict__ tile_excl) {
    const uint32_t gw = (blockIdx.x * kThreads + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t base = gw * kVTile;
    if (base >= n) return;
    VScan carry = tile_excl[gw];
#pragma unroll 1
    for (int r = 0; r < kVTile / 32; ++r) {
        const uint32_t i = base + r * 32 + lane;
        const VScan v = i < n ? a[i] : vs_identity();
        const VScan inc = vs_warp_scan(v, lane);
        VScan ex = vs_shfl_up(inc, 1);
        if (lane == 0) ex = vs_identity();
        if (i < n) a[i] = vs_combine(carry, ex);
        carry = vs_combine(carry, vs_shfl(inc, 31));
    }
}

// The element pass keeps the round-based traversal (lane = consecutive event): its 64 B of output per event leave
// the warp as whole contiguous lines.  (A lane-serial version — as in k_vs_reduce_elems — halves the shuffles but
// writes 32 separate sectors per store instruction: 279 us against 166 us per 10M events.)
__global__ void __launch_bounds__(kThreads)
k_vs_apply_elems(VCols c, const VScan* __restrict__ tile_excl, igs_prefix_rec* __restrict__ pre,
                 igs_tail_rec* __restrict__ tail) {
    const uint32_t gw = (blockIdx.x * kThreads + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t base = gw * kVTile;
    if (base >= c.n) return;
    VScan carry = tile_excl[gw];
#pragma unroll 1
    for (int r = 0; r < kVTile / 32; ++r) {
        const uint32_t i = base + r * 32 + lane;
        const VScan v = i < c.n ? vs_element(c, i) : vs_identity();
        const VScan inc = vs_warp_scan(v, lane);
        VScan ex = vs_shfl_up(inc, 1);
        if (lane == 0) ex = vs_identity();
        if (i < c.n) {
            const VScan before = vs_combine(carry, ex);          // everything ordered before event i
            const VScan upto = vs_combine(carry, inc);           // ... including event i
            igs_prefix_rec p;
            p.dcnt = before.d; p.cnt[0] = before.c0; p.cnt[1] = before.c1; p.pad = 0;
            p.sum[0] = before.s0; p.sum[1] = before.s1;
            pre[i] = p;
            igs_tail_rec tl;
            tl.chain_sum = upto.chain; tl.sess_start = upto.st;
            tl.dev_ts = tl.ip_ts = 0; tl.dev_cnt = tl.ip_cnt = 0; tl.pad[0] = tl.pad[1] = 0;   // step 5 fills these
            tail[i] = tl;
            if (i == c.n - 1) {                                  // pre[n]: totals over the whole history
                igs_prefix_rec q;
                q.dcnt = upto.d; q.cnt[0] = upto.c0; q.cnt[1] = upto.c1; q.pad = 0;
                q.sum[0] = upto.s0; q.sum[1] = upto.s1;
                pre[c.n] = q;
            }
        }
        carry = vs_combine(carry, vs_shfl(inc, 31));
    }
}

// ------------------------------------------------------------------------------------------------
// query kernel
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
k_velocity(igs_history_view ix, const igs_query* __restrict__ q, unsigned long long n,
           igs_velocity* __restrict__ out) {
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 qq = __ldg(reinterpret_cast<const uint4*>(q + i));
    igs_velocity_a16 v;
    igs_velocity_lookup(ix, qq.x, qq.y, qq.z, &v.v);
    uint4* dst = reinterpret_cast<uint4*>(out + i);
#pragma unroll
    for (int k = 0; k < (int)(sizeof(igs_velocity) / 16); ++k) dst[k] = v.q[k];
}

__global__ void __launch_bounds__(kThreads)
k_uniques(igs_history_view ix, const igs_query* __restrict__ q, unsigned long long n,
          igs_uniques* __restrict__ out) {
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 qq = __ldg(reinterpret_cast<const uint4*>(q + i));
    igs_velocity_a16 v;
    igs_uniques u;
    igs_velocity_lookup(ix, qq.x, qq.y, qq.z, &v.v, &u);
    out[i] = u;
}

// ------------------------------------------------------------------------------------------------
// identifier columns (step 5)
// ------------------------------------------------------------------------------------------------
constexpr int kIdKeyBits = 32;       // 4 radix passes of 8 bits

// id_s[p] = identifier of sorted event p; flag[p] = 1 iff it has one
__global__ void k_id_gather(const uint64_t* __restrict__ raw, const uint32_t* __restrict__ seq, uint32_t n,
                            uint64_t* __restrict__ id_s, uint32_t* __restrict__ flag) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const uint64_t v = __ldg(raw + __ldg(seq + p));
    id_s[p] = v;
    flag[p] = v != 0;
}

// compacted sequence k = 0..m-1 of the events that carry an identifier (fex = exclusive scan of flag)
__global__ void k_id_compact(const uint64_t* __restrict__ id_s, const uint32_t* __restrict__ fex, uint32_t n,
                             uint32_t* __restrict__ cpos, uint64_t* __restrict__ cid) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const uint64_t v = __ldg(id_s + p);
    if (v) {
        const uint32_t k = __ldg(fex + p);
        cpos[k] = p;
        cid[k] = v;
    }
}

// chain heads: the key did not exist when event k arrived (new account, or the previous PFADD was more than
// 24h earlier and the key had expired, :144/:151)
__global__ void k_id_heads(const uint32_t* __restrict__ cpos, const uint32_t* __restrict__ acct,
                           const uint32_t* __restrict__ ts, uint32_t m, uint8_t* __restrict__ head,
                           uint32_t* __restrict__ hsum) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= m) return;
    bool h = k == 0;
    if (!h) {
        const uint32_t p = __ldg(cpos + k), q = __ldg(cpos + k - 1);
        h = __ldg(acct + p) != __ldg(acct + q) ||
            (long long)__ldg(ts + p) - (long long)__ldg(ts + q) > 86400;
    }
    head[k] = h;
    hsum[k] = h;
}

// hsum holds the exclusive scan of the head flags on entry and the chain number of k on exit
__global__ void k_id_chain(const uint8_t* __restrict__ head, uint32_t* __restrict__ hsum,
                           const uint64_t* __restrict__ cid, uint32_t m, uint32_t* __restrict__ chain_head,
                           uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= m) return;
    const uint32_t h = __ldg(head + k);
    const uint32_t g = hsum[k] + h - 1;
    hsum[k] = g;
    if (h) chain_head[g] = k;
    keys[k] = igs_idset_mix(__ldg(cid + k)) >> (64 - kIdKeyBits);
    vals[k] = k;
}

// After the stable sort by hash, the entries of one hash value are in event order, so the events of one
// chain are adjacent; the first earlier entry of the same chain with the same 64-bit id — normally the
// entry right before — is a previous occurrence.  (Different ids with equal hashes only lengthen the walk.)
__global__ void k_id_new(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                         const uint32_t* __restrict__ chain, const uint64_t* __restrict__ cid, uint32_t m,
                         uint8_t* __restrict__ isnew8, uint32_t* __restrict__ isnew32) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= m) return;
    const uint32_t k = __ldg(vals + r);
    const uint64_t key = __ldg(keys + r), id = __ldg(cid + k);
    const uint32_t g = __ldg(chain + k);
    bool nw = true;
    for (uint32_t rr = r; rr-- > 0;) {
        if (__ldg(keys + rr) != key) break;
        const uint32_t k2 = __ldg(vals + rr);
        if (__ldg(chain + k2) != g) break;
        if (__ldg(cid + k2) == id) { nw = false; break; }
    }
    isnew8[k] = nw;
    isnew32[k] = nw;
}

// state of the key as of every sorted event p (with or without an identifier of its own)
template <int kCol>
__global__ void k_id_expand(const uint64_t* __restrict__ id_s, const uint32_t* __restrict__ fex,
                            const uint32_t* __restrict__ cpos, const uint32_t* __restrict__ chain,
                            const uint32_t* __restrict__ chain_head, const uint32_t* __restrict__ nex,
                            const uint8_t* __restrict__ isnew8, const uint32_t* __restrict__ acct,
                            const uint32_t* __restrict__ ts, uint32_t n, igs_tail_rec* __restrict__ tail) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const uint32_t kk = __ldg(fex + p) + (__ldg(id_s + p) != 0);   // identifier events at or before p
    uint32_t cnt = 0, t = 0;
    if (kk) {
        const uint32_t k = kk - 1, q = __ldg(cpos + k);
        if (__ldg(acct + q) == __ldg(acct + p)) {
            const uint32_t h = __ldg(chain_head + __ldg(chain + k));
            cnt = __ldg(nex + k) + __ldg(isnew8 + k) - __ldg(nex + h);   // first occurrences in [h, k]
            t = __ldg(ts + q);
        }
    }
    const uint16_t c16 = (uint16_t)min(cnt, 65535u);
    if (kCol == 0) { tail[p].dev_ts = t; tail[p].dev_cnt = c16; }
    else           { tail[p].ip_ts = t;  tail[p].ip_cnt = c16; }
}

// ------------------------------------------------------------------------------------------------
// eviction (igs_evict_events)
// ------------------------------------------------------------------------------------------------
// account of sorted position j = the segment that contains it
__device__ __forceinline__ uint32_t account_of(const uint32_t* __restrict__ seg, uint32_t n_accounts, uint32_t j) {
    uint32_t lo = 0, hi = n_accounts;                 // first a with seg[a + 1] > j
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if (__ldg(seg + mid + 1) <= j) lo = mid + 1; else hi = mid;
    }
    return lo;
}
// keep_from[a] = first sorted position of account a that can still influence a query at t >= now: the start of
// the trailing chain of events with consecutive gaps <= 86400 (the longest key life among the chains: tx_sum 1h,
// session 30m, identifier keys 24h; windows reach back 24h), or the end of the segment when even last_tx (7d,
// the longest TTL of all, redis_store.go:155) has expired.
// A break after which the account's later events all lie in the future of `now` (future-dated events, a caller
// whose clock is behind the ingest clock) must not drop what precedes it: queries at now <= t < ts(break event)
// still see those events.  Only breaks whose later event has ts <= now count.
__global__ void k_ev_breaks(const uint32_t* __restrict__ seg, const uint32_t* __restrict__ ts, uint32_t n,
                            uint32_t n_accounts, uint32_t now, uint32_t* __restrict__ keep_from) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint32_t t = __ldg(ts + j);
    if (t > now) return;
    const bool maybe = j == 0 || t < __ldg(ts + j - 1) || (long long)t - (long long)__ldg(ts + j - 1) > 86400;
    if (!maybe) return;                               // same account, gap <= 24h: not a break
    const uint32_t a = account_of(seg, n_accounts, j);
    const bool brk = j == __ldg(seg + a) || (long long)t - (long long)__ldg(ts + j - 1) > 86400;
    if (brk) atomicMax(keep_from + a, j);
}
__global__ void k_ev_dead(const uint32_t* __restrict__ seg, const uint32_t* __restrict__ ts, uint32_t n_accounts,
                          uint32_t now, uint32_t* __restrict__ keep_from) {
    const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n_accounts) return;
    const uint32_t s = __ldg(seg + a), e = __ldg(seg + a + 1);
    if (e == s) return;
    if ((long long)now - (long long)__ldg(ts + e - 1) > 7 * 86400) keep_from[a] = e;
}
// keep flag per ARRIVAL index (the raw columns are in arrival order)
__global__ void k_ev_keep(const uint32_t* __restrict__ seg, const uint32_t* __restrict__ seq, uint32_t n,
                          uint32_t n_accounts, const uint32_t* __restrict__ keep_from, uint32_t* __restrict__ keep) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint32_t a = account_of(seg, n_accounts, j);
    keep[__ldg(seq + j)] = j >= __ldg(keep_from + a);
}
__global__ void k_save_last_u32(const uint32_t* __restrict__ flag, uint32_t n, uint32_t* __restrict__ out) {
    if (blockIdx.x == 0 && threadIdx.x == 0) *out = flag[n - 1];
}
template <typename T>
__global__ void k_ev_compact(const T* __restrict__ in, const uint32_t* __restrict__ keep_ex,
                             const uint32_t* __restrict__ last_keep, uint32_t n, T* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t ex = __ldg(keep_ex + i);
    const bool kept = i + 1 < n ? __ldg(keep_ex + i + 1) != ex : *last_keep != 0;
    if (kept) out[ex] = in[i];
}

int bits_for(uint64_t v) {   // number of bits needed to represent v
    int b = 0;
    while (v) { ++b; v >>= 1; }
    return b;
}

void free_build_pool(igs_engine* e) {
    for (igs_buf& b : e->build_pool)
        if (b.p) cudaFree(b.p);
    e->build_pool.clear();
}

igs_index_view view_of(const igs_event_index& x, bool valid) {
    igs_index_view v;
    v.seg = x.seg; v.ts = x.ts; v.seq = x.seq; v.pre = x.pre; v.tail = x.tail;
    v.n_accounts = valid ? x.n_accounts : 0;
    v.n = valid ? (uint32_t)x.n : 0;
    return v;
}
igs_index_view no_view() {
    igs_index_view v;
    v.seg = nullptr; v.ts = nullptr; v.seq = nullptr; v.pre = nullptr; v.tail = nullptr;
    v.n_accounts = 0; v.n = 0;
    return v;
}

void release_index_arrays(igs_engine* e, igs_event_index& x) {
    if (x.seg) cudaFree(x.seg);
    if (x.ts) cudaFree(x.ts);
    if (x.seq) cudaFree(x.seq);
    if (x.pre) cudaFree(x.pre);
    if (x.tail) cudaFree(x.tail);
    for (cudaEvent_t ev : x.readers) e->free_events.push_back(ev);   // cudaFree has waited for the device
    x = igs_event_index();
}

// Makes `st` wait for every kernel that was launched while x's arrays were the published ones.  Caller holds view_mu.
void wait_for_readers(igs_engine* e, igs_event_index& x, cudaStream_t st) {
    for (cudaEvent_t ev : x.readers) {
        cudaStreamWaitEvent(st, ev, 0);
        e->free_events.push_back(ev);
    }
    x.readers.clear();
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// published view + reader bookkeeping
// ------------------------------------------------------------------------------------------------
igs_history_reader::igs_history_reader(igs_engine* eng) : e(eng), locked(true) {
    e->view_mu.lock();
    view.base = view_of(e->ev, e->has_events);
    view.delta = e->has_events && e->dv.n != 0 ? view_of(e->dv, true) : no_view();
}
void igs_history_reader::done(cudaStream_t st) {
    if (!locked) return;
    igs_event_index* levels[2] = {&e->ev, &e->dv};
    for (igs_event_index* x : levels) {
        if (x->n == 0 || !e->has_events) continue;
        if (x->readers.size() >= 64) {               // drop the ones that have completed
            size_t k = 0;
            for (cudaEvent_t ev : x->readers) {
                if (cudaEventQuery(ev) == cudaSuccess) e->free_events.push_back(ev);
                else x->readers[k++] = ev;
            }
            x->readers.resize(k);
            cudaGetLastError();                      // cudaErrorNotReady is not an error
        }
        cudaEvent_t ev = nullptr;
        if (!e->free_events.empty()) { ev = e->free_events.back(); e->free_events.pop_back(); }
        else if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); continue; }
        if (cudaEventRecord(ev, st) == cudaSuccess) x->readers.push_back(ev);
        else { cudaGetLastError(); e->free_events.push_back(ev); }
    }
    locked = false;
    e->view_mu.unlock();
}
igs_history_reader::~igs_history_reader() {
    if (locked) e->view_mu.unlock();
}

igs_history_view igs_history_view_of(igs_engine* e) {
    std::lock_guard<std::mutex> lk(e->view_mu);
    igs_history_view v;
    v.base = view_of(e->ev, e->has_events);
    v.delta = e->has_events && e->dv.n != 0 ? view_of(e->dv, true) : no_view();
    return v;
}

static void free_raw_events(igs_engine* e) {
    if (e->raw_account) cudaFree(e->raw_account);
    if (e->raw_ts) cudaFree(e->raw_ts);
    if (e->raw_amount) cudaFree(e->raw_amount);
    if (e->raw_type) cudaFree(e->raw_type);
    if (e->raw_device) cudaFree(e->raw_device);
    if (e->raw_ip) cudaFree(e->raw_ip);
    e->raw_account = e->raw_ts = nullptr; e->raw_amount = nullptr; e->raw_type = nullptr;
    e->raw_device = e->raw_ip = nullptr;
    e->has_event_ids = false;
    e->n_raw = e->cap_raw = e->base_n = 0;
}

void igs_free_event_index(igs_engine* e) {
    {
        std::lock_guard<std::mutex> lk(e->view_mu);
        e->has_events = false;
        for (igs_event_index* x : {&e->ev, &e->ev_spare, &e->dv, &e->dv_spare}) release_index_arrays(e, *x);
    }
    free_raw_events(e);
    free_build_pool(e);
}

void igs_free_history_events(igs_engine* e) {       // engine teardown: the pooled events
    for (cudaEvent_t ev : e->free_events) cudaEventDestroy(ev);
    e->free_events.clear();
}

// Appends n events to the device-resident raw columns (capacity grows by 1.5x).
// (the source columns may be host or device memory: cudaMemcpyDefault lets the runtime tell)
static int reserve_raw(igs_engine* e, uint64_t need, cudaStream_t st);
static int append_raw(igs_engine* e, const igs_event_cols* ev, const igs_event_ids* ids, uint64_t n,
                      cudaStream_t st) {
    const uint64_t need = e->n_raw + n;
    {
        const int rc = reserve_raw(e, need, st);
        if (rc != IGS_OK) return rc;
    }
    if (n) {
        IGS_CUDA(e, cudaMemcpyAsync(e->raw_account + e->n_raw, ev->account, n * 4, cudaMemcpyDefault, st));
        IGS_CUDA(e, cudaMemcpyAsync(e->raw_ts + e->n_raw, ev->ts, n * 4, cudaMemcpyDefault, st));
        IGS_CUDA(e, cudaMemcpyAsync(e->raw_amount + e->n_raw, ev->amount, n * 8, cudaMemcpyDefault, st));
        IGS_CUDA(e, cudaMemcpyAsync(e->raw_type + e->n_raw, ev->tx_type, n, cudaMemcpyDefault, st));
        if (e->has_event_ids) {       // a missing column = events without that identifier
            if (ids && ids->device)
                IGS_CUDA(e, cudaMemcpyAsync(e->raw_device + e->n_raw, ids->device, n * 8, cudaMemcpyDefault, st));
            else
                IGS_CUDA(e, cudaMemsetAsync(e->raw_device + e->n_raw, 0, n * 8, st));
            if (ids && ids->ip)
                IGS_CUDA(e, cudaMemcpyAsync(e->raw_ip + e->n_raw, ids->ip, n * 8, cudaMemcpyDefault, st));
            else
                IGS_CUDA(e, cudaMemsetAsync(e->raw_ip + e->n_raw, 0, n * 8, st));
        }
    }
    e->n_raw = need;
    return IGS_OK;
}

// Grows the raw columns to hold at least `need` events (capacity grows by 1.5x); the resident events are kept.
static int reserve_raw(igs_engine* e, uint64_t need, cudaStream_t st) {
    if (need <= e->cap_raw) return IGS_OK;
    const uint64_t cap = e->cap_raw == 0 ? need : need + need / 2 + 1024;
    {
        uint32_t *a = nullptr, *t = nullptr; int64_t* m = nullptr; uint8_t* y = nullptr;
        uint64_t *d = nullptr, *p = nullptr;
        {
            cudaError_t err = cudaMalloc(&a, cap * 4);
            if (err == cudaSuccess) err = cudaMalloc(&t, cap * 4);
            if (err == cudaSuccess) err = cudaMalloc(&m, cap * 8);
            if (err == cudaSuccess) err = cudaMalloc(&y, cap);
            if (err == cudaSuccess && e->has_event_ids) err = cudaMalloc(&d, cap * 8);
            if (err == cudaSuccess && e->has_event_ids) err = cudaMalloc(&p, cap * 8);
            if (err != cudaSuccess) {                 // give back what was obtained; the old columns are untouched
                for (void* q : {(void*)a, (void*)t, (void*)m, (void*)y, (void*)d, (void*)p})
                    if (q) cudaFree(q);
                igs_set_error(e, "growing the event columns to %llu events failed: %s", (unsigned long long)cap,
                              cudaGetErrorString(err));
                return err == cudaErrorMemoryAllocation ? IGS_ERR_NOMEM : IGS_ERR_CUDA;
            }
        }
        if (e->n_raw) {
            IGS_CUDA(e, cudaMemcpyAsync(a, e->raw_account, e->n_raw * 4, cudaMemcpyDeviceToDevice, st));
            IGS_CUDA(e, cudaMemcpyAsync(t, e->raw_ts, e->n_raw * 4, cudaMemcpyDeviceToDevice, st));
            IGS_CUDA(e, cudaMemcpyAsync(m, e->raw_amount, e->n_raw * 8, cudaMemcpyDeviceToDevice, st));
            IGS_CUDA(e, cudaMemcpyAsync(y, e->raw_type, e->n_raw, cudaMemcpyDeviceToDevice, st));
            if (e->has_event_ids) {
                IGS_CUDA(e, cudaMemcpyAsync(d, e->raw_device, e->n_raw * 8, cudaMemcpyDeviceToDevice, st));
                IGS_CUDA(e, cudaMemcpyAsync(p, e->raw_ip, e->n_raw * 8, cudaMemcpyDeviceToDevice, st));
            }
            IGS_CUDA(e, cudaStreamSynchronize(st));
        }
        if (e->raw_account) { cudaFree(e->raw_account); cudaFree(e->raw_ts); cudaFree(e->raw_amount); cudaFree(e->raw_type); }
        if (e->raw_device) { cudaFree(e->raw_device); cudaFree(e->raw_ip); }
        e->raw_account = a; e->raw_ts = t; e->raw_amount = m; e->raw_type = y;
        e->raw_device = d; e->raw_ip = p;
        e->cap_raw = cap;
    }
    return IGS_OK;
}

// Settles the distinct flags of the members of long (account, ts) groups (k_distinct's work list).  The table comes
// from the build pool like everything else and is requested on every build (one element when there is nothing to
// settle), so that the pool's request sequence stays fixed and rebuilds reuse the buffers.
static int resolve_big_groups(igs_engine* e, Scratch& sc, uint32_t n_work, const uint32_t* d_work, const uint32_t* d_acct,
                              const uint32_t* d_ts, const int64_t* d_amt, uint8_t* d_dflag, cudaStream_t st) {
    uint32_t cap = 1;
    if (n_work) {
        cap = 64;
        while (cap < 2 * (uint64_t)n_work && cap < 0x80000000u) cap <<= 1;
    }
    uint32_t *rep, *minpos;
    IGS_CUDA(e, sc.alloc(&rep, cap));
    IGS_CUDA(e, sc.alloc(&minpos, cap));
    if (!n_work) return IGS_OK;
    IGS_CUDA(e, cudaMemsetAsync(rep, 0xFF, (size_t)cap * 4, st));
    IGS_CUDA(e, cudaMemsetAsync(minpos, 0xFF, (size_t)cap * 4, st));
    const uint32_t nb = (n_work + kThreads - 1) / kThreads;
    k_dg_insert<<<nb, kThreads, 0, st>>>(d_work, n_work, d_acct, d_ts, d_amt, cap - 1, rep, minpos);
    IGS_LAUNCHED(e);
    k_dg_resolve<<<nb, kThreads, 0, st>>>(d_work, n_work, d_acct, d_ts, d_amt, cap - 1, rep, minpos, d_dflag);
    IGS_LAUNCHED(e);
    return IGS_OK;
}

struct IdScratch {
    uint64_t* keys[2]; uint32_t* vals[2];
    uint64_t *id_s, *cid;
    uint32_t *fex, *cpos, *chain, *chain_head, *nex, *hist, *scan_scratch;
    uint8_t *head, *isnew8;
};

// Step 5 for one identifier column: fills tail[p].{dev,ip}_{ts,cnt}.  d_acct / x.ts / x.seq are the sorted
// columns of the build in progress; `raw` is the whole raw column (x.seq holds global arrival indices).
template <int kCol>
static int build_id_column(igs_engine* e, igs_event_index& x, IdScratch& s, const uint64_t* raw, const uint32_t* d_acct,
                           uint32_t n, cudaStream_t st) {
    const uint32_t nb = (n + kThreads - 1) / kThreads;
    k_id_gather<<<nb, kThreads, 0, st>>>(raw, x.seq, n, s.id_s, s.fex);
    IGS_LAUNCHED(e);
    int rc = scan_u32(e, s.fex, n, s.scan_scratch, st);
    if (rc != IGS_OK) return rc;
    uint32_t last_ex = 0; uint64_t last_id = 0;
    IGS_CUDA(e, cudaMemcpyAsync(&last_ex, s.fex + (n - 1), 4, cudaMemcpyDeviceToHost, st));
    IGS_CUDA(e, cudaMemcpyAsync(&last_id, s.id_s + (n - 1), 8, cudaMemcpyDeviceToHost, st));
    IGS_CUDA(e, cudaStreamSynchronize(st));
    const uint32_t m = last_ex + (last_id != 0);       // events that carry this identifier
    if (m == 0) return IGS_OK;                         // the tail records already say "no key"
    const uint32_t mb = (m + kThreads - 1) / kThreads;
    k_id_compact<<<nb, kThreads, 0, st>>>(s.id_s, s.fex, n, s.cpos, s.cid);
    IGS_LAUNCHED(e);
    k_id_heads<<<mb, kThreads, 0, st>>>(s.cpos, d_acct, x.ts, m, s.head, s.chain);
    IGS_LAUNCHED(e);
    rc = scan_u32(e, s.chain, m, s.scan_scratch, st);
    if (rc != IGS_OK) return rc;
    k_id_chain<<<mb, kThreads, 0, st>>>(s.head, s.chain, s.cid, m, s.chain_head, s.keys[0], s.vals[0]);
    IGS_LAUNCHED(e);
    int cur = 0;
    rc = radix_sort_pairs(e, s.keys, s.vals, m, kIdKeyBits / kSortMaxBits, kSortMaxBits, s.hist, s.scan_scratch, st, &cur);
    if (rc != IGS_OK) return rc;
    k_id_new<<<mb, kThreads, 0, st>>>(s.keys[cur], s.vals[cur], s.chain, s.cid, m, s.isnew8, s.nex);
    IGS_LAUNCHED(e);
    rc = scan_u32(e, s.nex, m, s.scan_scratch, st);
    if (rc != IGS_OK) return rc;
    k_id_expand<kCol><<<nb, kThreads, 0, st>>>(s.id_s, s.fex, s.cpos, s.chain, s.chain_head, s.nex, s.isnew8, d_acct,
                                               x.ts, n, x.tail);
    IGS_LAUNCHED(e);
    return IGS_OK;
}

// One index build: raw events [lo, hi) -> the arrays of *dst (not published; the caller swaps them in).
//   base      carry-in for a delta build (base.n != 0): chain columns continue the base's, the distinct flags know
//             the base's last (account, ts) group, a delta event older than its account's last base event is late
//   n_before  arrival indices >= n_before are the batch being appended: late[1] counts the late pairs among them
//             (late[0]: all late pairs of the range)
// Runs on the ingest stream and returns after its kernels have completed.
struct BuildSpec {
    uint64_t lo = 0, hi = 0;
    uint32_t n_accounts = 0;
    igs_event_index* dst = nullptr;
    igs_index_view base;
    uint64_t n_before = 0xFFFFFFFFull;
    bool assume_time_ordered = false;   // delta builds: sort by account only, no key-range pass (see build_index)
    uint64_t reserve = 0;               // size allocations for at least this many events
};

// (Re)allocates the arrays of an unpublished index so that they hold `events` events of `n_accounts` accounts, with
// headroom so that appends rarely come back here.
static int ensure_index_capacity(igs_engine* e, igs_event_index& x, uint64_t events, uint32_t n_accounts) {
    if (events + 1 <= x.cap_events && (uint64_t)n_accounts + 1 <= x.cap_accounts) return IGS_OK;
    // first allocation: what is asked for; growth: half as much again, so that a history that grows by an eighth
    // per compaction comes back here every few compactions, not every time
    const uint64_t ce = x.cap_events == 0 ? events + 1 : events + 1 + events / 2;
    const uint64_t ca = std::max<uint64_t>((uint64_t)n_accounts + 1, x.cap_accounts);
    {
        std::lock_guard<std::mutex> lk(e->view_mu);
        release_index_arrays(e, x);
    }
    IGS_CUDA(e, cudaMalloc(&x.seg, ca * sizeof(uint32_t)));
    IGS_CUDA(e, cudaMalloc(&x.ts, ce * sizeof(uint32_t)));
    IGS_CUDA(e, cudaMalloc(&x.seq, ce * sizeof(uint32_t)));
    IGS_CUDA(e, cudaMalloc(&x.pre, ce * sizeof(igs_prefix_rec)));
    IGS_CUDA(e, cudaMalloc(&x.tail, ce * sizeof(igs_tail_rec)));
    x.cap_events = ce;
    x.cap_accounts = ca;
    return IGS_OK;
}

static int build_index(igs_engine* e, const BuildSpec& spec, uint32_t late[2]) {
    const uint32_t n = (uint32_t)(spec.hi - spec.lo);
    const uint32_t n_accounts = spec.n_accounts;
    cudaStream_t st = e->ingest_stream;
    igs_event_index& x = *spec.dst;
    Scratch sc(e->build_pool);
    late[0] = late[1] = 0;
    // sizes of allocations: at least what the caller expects the range to grow to (a delta is sized for its
    // compaction threshold once, instead of being re-allocated every few appends)
    const uint32_t nr = (uint32_t)std::max<uint64_t>(n, std::min<uint64_t>(spec.reserve, 0xFFFFFFF0ull));
    {
        const int rc = ensure_index_capacity(e, x, nr, n_accounts);
        if (rc != IGS_OK) return rc;
    }
    x.n = n;
    x.n_accounts = n_accounts;
    IGS_CUDA(e, cudaMemsetAsync(x.seg, 0, ((size_t)n_accounts + 1) * sizeof(uint32_t), st));
    IGS_CUDA(e, cudaMemsetAsync(x.pre, 0, sizeof(igs_prefix_rec), st));
    if (n == 0) {
        IGS_CUDA(e, cudaStreamSynchronize(st));
        return IGS_OK;
    }

    // ---- the raw columns are already on the device -----------------------------------------------
    const uint32_t* d_acct_in = e->raw_account + spec.lo;
    const uint32_t* d_ts_in = e->raw_ts + spec.lo;
    const int64_t* d_amt_in = e->raw_amount + spec.lo;
    const uint8_t* d_type_in = e->raw_type + spec.lo;
    uint32_t* d_mm;
    IGS_CUDA(e, sc.alloc(&d_mm, 8));            // [0..3] key range, [4..5] late-event counters, [6] bad account
    const uint32_t mm_init[8] = {0xFFFFFFFFu, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    IGS_CUDA(e, cudaMemcpyAsync(d_mm, mm_init, sizeof(mm_init), cudaMemcpyHostToDevice, st));
    const uint32_t nb_el = (n + kThreads - 1) / kThreads;

    // ---- key range ------------------------------------------------------------------------------
    // A delta build skips this pass and its read-back: its events must be in time order per account anyway (a late
    // event gets the batch refused, or everything re-sorted into a new base), so a stable sort by account alone is
    // the right order, and the account range is that of the table.
    uint32_t ts_min = 0, acct_max = n_accounts ? n_accounts - 1 : 0;
    bool time_ordered = true;
    uint32_t ts_span = 0;
    if (!spec.assume_time_ordered) {
        k_minmax<<<min(nb_el, (uint32_t)e->sm_count * 8), kThreads, 0, st>>>(d_acct_in, d_ts_in, n, d_mm);
        IGS_LAUNCHED(e);
        uint32_t mm[4];
        IGS_CUDA(e, cudaMemcpyAsync(mm, d_mm, sizeof(mm), cudaMemcpyDeviceToHost, st));
        IGS_CUDA(e, cudaStreamSynchronize(st));
        if (mm[2] >= n_accounts) {
            igs_set_error(e, "event account index %u >= n_accounts %u", mm[2], n_accounts);
            return IGS_ERR_INVALID;
        }
        ts_min = mm[0];
        ts_span = mm[1] - mm[0];
        acct_max = mm[2];
        // Events that arrive in time order (a log, a queue) only need a stable sort by account.
        time_ordered = mm[3] == 0;
    }
    const int ts_bits = time_ordered ? 0 : (bits_for(ts_span) > 0 ? bits_for(ts_span) : 1);
    const int acct_bits = bits_for(acct_max) > 0 ? bits_for(acct_max) : 1;
    const int seq_bits = bits_for(n - 1) > 0 ? bits_for(n - 1) : 1;
    static const bool force_pairs = getenv("IGS_SORT_PAIRS") && atoi(getenv("IGS_SORT_PAIRS")) != 0;   // test knob: old path
    const bool packed = !force_pairs && n < (1u << 30) && seq_bits + 2 + ts_bits + acct_bits <= 64;

    uint32_t* d_acct; int64_t* d_amt; uint8_t *d_type, *d_dflag;
    uint64_t* d_keys[2]; uint32_t* d_vals[2];
    IGS_CUDA(e, sc.alloc(&d_keys[0], nr));
    IGS_CUDA(e, sc.alloc(&d_keys[1], nr));
    const bool need_vals = !packed || e->has_event_ids;               // the identifier columns sort (hash, position) pairs
    IGS_CUDA(e, sc.alloc(&d_vals[0], need_vals ? nr : 1));
    IGS_CUDA(e, sc.alloc(&d_vals[1], need_vals ? nr : 1));
    IGS_CUDA(e, sc.alloc(&d_acct, nr));
    IGS_CUDA(e, sc.alloc(&d_amt, nr));
    IGS_CUDA(e, sc.alloc(&d_type, nr));
    IGS_CUDA(e, sc.alloc(&d_dflag, nr));
    const uint32_t nb_sort = (n + kSortTile - 1) / kSortTile;
    if (packed) {
        // ---- one-sweep radix sort of packed keys (onesweep.cuh) ---------------------------------------
        KeyLayout lay;
        lay.seq_bits = seq_bits; lay.ts_bits = ts_bits; lay.acct_shift = seq_bits + 2 + ts_bits; lay.ts_min = ts_min;
        const OsPlan plan = os_plan(seq_bits + 2, seq_bits + 2 + ts_bits + acct_bits);
        uint32_t* d_os;
        IGS_CUDA(e, sc.alloc(&d_os, OsScratch::words(nr, kOsMaxPasses)));
        OsScratch osc;
        int rc = os_prepare(e, &osc, d_os, n, plan, st);
        if (rc != IGS_OK) return rc;
        IGS_CUDA(e, cudaFuncSetAttribute(k_make_keys_hist, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)os_hist_smem_bytes(plan.passes)));
        k_make_keys_hist<<<min((n + kOsThreads - 1) / kOsThreads, (uint32_t)e->sm_count * 8), kOsThreads,
                           os_hist_smem_bytes(plan.passes), st>>>(d_acct_in, d_ts_in, d_type_in, n, n_accounts, lay, plan, osc,
                                                                  d_keys[0], d_mm + 6);
        IGS_LAUNCHED(e);
        int cur = 0;
        rc = os_sort_passes(e, d_keys, n, plan, osc, st, &cur);
        if (rc != IGS_OK) return rc;
        IGS_CUDA(e, cudaMemcpyAsync(d_mm + 7, osc.error(), 4, cudaMemcpyDeviceToDevice, st));
        k_unpack_gather_packed<<<nb_el, kThreads, 0, st>>>(d_keys[cur], d_ts_in, d_amt_in, n, lay, (uint32_t)spec.lo, d_acct,
                                                           x.ts, x.seq, d_amt, d_type);
        IGS_LAUNCHED(e);
    } else {
        // ---- (key, value) pairs: histories whose account / time / arrival bits do not fit one word --------
        if (spec.assume_time_ordered) {
            igs_set_error(e, "internal: a delta build needs packed sort keys");
            return IGS_ERR_INVALID;
        }
        const int key_bits = ts_bits + acct_bits;
        const int passes = (key_bits + kSortMaxBits - 1) / kSortMaxBits;
        const int digit_bits = (key_bits + passes - 1) / passes;      // <= 8
        const uint32_t bins = 1u << digit_bits;
        const uint32_t hist_elems = bins * nb_sort;
        uint32_t *d_hist, *d_scan_scratch;
        IGS_CUDA(e, sc.alloc(&d_hist, hist_elems));
        IGS_CUDA(e, sc.alloc(&d_scan_scratch, scan_u32_scratch_elems(hist_elems)));
        k_make_keys<<<nb_el, kThreads, 0, st>>>(d_acct_in, d_ts_in, n, ts_min, ts_bits, d_keys[0], d_vals[0]);
        IGS_LAUNCHED(e);
        int cur = 0;
        const int rc = radix_sort_pairs(e, d_keys, d_vals, n, passes, digit_bits, d_hist, d_scan_scratch, st, &cur);
        if (rc != IGS_OK) return rc;
        k_unpack_gather<<<nb_el, kThreads, 0, st>>>(d_keys[cur], d_vals[cur], d_ts_in, d_amt_in, d_type_in, n, ts_min, ts_bits,
                                                    (uint32_t)spec.lo, d_acct, x.ts, x.seq, d_amt, d_type);
        IGS_LAUNCHED(e);
    }

    // ---- segments, group heads, late events, distinct flags -----------------------------------------
    uint8_t* d_head;
    IGS_CUDA(e, sc.alloc(&d_head, nr));
    uint32_t* d_work;                                                  // members of long (account, ts) groups
    IGS_CUDA(e, sc.alloc(&d_work, nr));
    k_segments_heads<<<min(nb_el, (uint32_t)e->sm_count * 8), kThreads, 0, st>>>(d_acct, x.ts, x.seq, n, n_accounts, (uint32_t)spec.n_before, spec.base,
                                                 x.seg, d_head, d_mm + 4);
    IGS_LAUNCHED(e);
    IGS_CUDA(e, cudaMemsetAsync(d_mm, 0, 4, st));                     // d_mm[0] is free again: unresolved counter
    k_distinct<<<nb_el, kThreads, 0, st>>>(d_head, d_amt, n, d_dflag, d_work, d_mm);
    IGS_LAUNCHED(e);
    {
        uint32_t n_unres = 0, flags[4] = {0, 0, 0, 0};           // late[0], late[1], bad account, look-back gave up
        IGS_CUDA(e, cudaMemcpyAsync(&n_unres, d_mm, 4, cudaMemcpyDeviceToHost, st));
        IGS_CUDA(e, cudaMemcpyAsync(flags, d_mm + 4, 16, cudaMemcpyDeviceToHost, st));
        IGS_CUDA(e, cudaStreamSynchronize(st));
        late[0] = flags[0]; late[1] = flags[1];
        if (flags[2] != 0) {
            igs_set_error(e, "event account index %u >= n_accounts %u", flags[2], n_accounts);
            return IGS_ERR_INVALID;
        }
        if (flags[3] != 0) {
            igs_set_error(e, "internal: the radix sort's look-back gave up waiting for an earlier tile");
            return IGS_ERR_CUDA;
        }
        if (spec.assume_time_ordered && late[0] != 0) return IGS_OK;   // not in time order after all: the caller decides
        if (late[1] != 0 && e->late_policy == IGS_LATE_REJECT) return IGS_OK;   // the caller drops the batch: stop here
        const int rc = resolve_big_groups(e, sc, n_unres, d_work, d_acct, x.ts, d_amt, d_dflag, st);
        if (rc != IGS_OK) return rc;
    }
    if (spec.base.n != 0) {
        k_delta_boundary_dedupe<<<nb_el, kThreads, 0, st>>>(d_acct, x.ts, d_amt, x.seg, n, spec.base, e->raw_amount, d_dflag);
        IGS_LAUNCHED(e);
    }

    // ---- fused scan: reduce up two levels, scan the top, apply down ---------------------------------
    const uint32_t n1 = (n + kVTile - 1) / kVTile;            // level-0 aggregates (one per 256 events)
    const uint32_t n2 = (n1 + kVTile - 1) / kVTile;           // level-1 aggregates (one per 65 536 events)
    VScan *d_agg1, *d_agg2;
    IGS_CUDA(e, sc.alloc(&d_agg1, (nr + kVTile - 1) / kVTile));
    IGS_CUDA(e, sc.alloc(&d_agg2, ((nr + kVTile - 1) / kVTile + kVTile - 1) / kVTile));
    VCols cols{d_acct, x.ts, d_amt, d_type, d_dflag, n, spec.base};
    const uint32_t warps_per_block = kThreads / 32;
    const uint32_t g1 = (n1 + warps_per_block - 1) / warps_per_block, g2 = (n2 + warps_per_block - 1) / warps_per_block;
    k_vs_reduce_elems<<<g1, kThreads, 0, st>>>(cols, d_agg1);
    IGS_LAUNCHED(e);
    k_vs_reduce<AggSrc><<<g2, kThreads, 0, st>>>(AggSrc{d_agg1, n1}, d_agg2);
    IGS_LAUNCHED(e);
    k_vs_scan_small<<<1, 32, 0, st>>>(d_agg2, n2);
    IGS_LAUNCHED(e);
    k_vs_apply_aggs<<<g2, kThreads, 0, st>>>(d_agg1, n1, d_agg2);
    IGS_LAUNCHED(e);
    k_vs_apply_elems<<<g1, kThreads, 0, st>>>(cols, d_agg1, x.pre, x.tail);
    IGS_LAUNCHED(e);

    // ---- identifier columns ---------------------------------------------------------------------
    if (e->has_event_ids) {
        IdScratch is;
        is.keys[0] = d_keys[0]; is.keys[1] = d_keys[1]; is.vals[0] = d_vals[0]; is.vals[1] = d_vals[1];   // the sort is done with them
        IGS_CUDA(e, sc.alloc(&is.id_s, n));
        IGS_CUDA(e, sc.alloc(&is.fex, n));
        IGS_CUDA(e, sc.alloc(&is.cpos, n));
        IGS_CUDA(e, sc.alloc(&is.cid, n));
        IGS_CUDA(e, sc.alloc(&is.head, n));
        IGS_CUDA(e, sc.alloc(&is.chain, n));
        IGS_CUDA(e, sc.alloc(&is.chain_head, n));
        IGS_CUDA(e, sc.alloc(&is.isnew8, n));
        IGS_CUDA(e, sc.alloc(&is.nex, n));
        IGS_CUDA(e, sc.alloc(&is.hist, (size_t)kSortBins * nb_sort));
        IGS_CUDA(e, sc.alloc(&is.scan_scratch, scan_u32_scratch_elems(std::max<uint32_t>((uint32_t)kSortBins * nb_sort, n))));
        int rc = build_id_column<0>(e, x, is, e->raw_device, d_acct, n, st);
        if (rc != IGS_OK) return rc;
        rc = build_id_column<1>(e, x, is, e->raw_ip, d_acct, n, st);
        if (rc != IGS_OK) return rc;
    }
    IGS_CUDA(e, cudaStreamSynchronize(st));
    return IGS_OK;
}

static int reject_late(igs_engine* e, uint32_t n_late) {
    igs_set_error(e, "the batch holds %u late event(s): an event older than one of the same account that is already "
                  "resident or earlier in the batch (policy IGS_LATE_REJECT; see igs_set_late_policy)", n_late);
    return IGS_ERR_INVALID;
}

// (Re)builds the base over every resident raw event and publishes it; the delta becomes empty.  A bulk load, an
// eviction, a compaction, an append that cannot go through the delta.  On any failure — a bad account index, a
// rejected late batch — nothing is published and the caller restores n_raw.
static int rebuild_base(igs_engine* e, uint32_t n_accounts, uint64_t n_before) {
    {
        std::lock_guard<std::mutex> lk(e->view_mu);
        wait_for_readers(e, e->ev_spare, e->ingest_stream);
    }
    BuildSpec spec;
    spec.lo = 0; spec.hi = e->n_raw; spec.n_accounts = n_accounts; spec.dst = &e->ev_spare;
    spec.base = no_view(); spec.n_before = n_before;
    uint32_t late[2];
    const int rc = build_index(e, spec, late);
    if (rc != IGS_OK) return rc;
    if (late[1] != 0 && e->late_policy == IGS_LATE_REJECT) return reject_late(e, late[1]);
    std::lock_guard<std::mutex> lk(e->view_mu);
    std::swap(e->ev, e->ev_spare);
    e->dv.n = 0;                         // its arrays stay until the next delta build has waited for their readers
    e->base_n = e->n_raw;
    e->late_events = late[0];
    e->has_events = true;
    return IGS_OK;
}

// Appended events sit in a delta index until it outgrows max(delta_min, base / 8); then base and delta are rebuilt
// into one base (amortised: a delta rebuild costs what the delta holds, a compaction what everything holds).

int igs_append_event_columns(igs_engine* e, const igs_event_cols* ev, const igs_event_ids* ids, uint64_t n) {
    if (!e->has_events) {
        igs_set_error(e, "igs_append_events needs a loaded history (igs_load_events, possibly with n = 0)");
        return IGS_ERR_STATE;
    }
    if (ids && !e->has_event_ids) {
        igs_set_error(e, "the loaded history carries no identifiers (load it with igs_load_events_ids)");
        return IGS_ERR_STATE;
    }
    if (e->n_raw + n >= 0xFFFFFFFFull) {
        igs_set_error(e, "at most 2^32-2 events per shard");
        return IGS_ERR_INVALID;
    }
    if (n == 0) return IGS_OK;
    const uint64_t before = e->n_raw;
    const uint32_t n_accounts = e->ev.n_accounts;
    int rc = append_raw(e, ev, ids, n, e->ingest_stream);
    if (rc != IGS_OK) { e->n_raw = before; return rc; }
    // Histories with identifier columns still rebuild everything: a delta would have to know which identifiers
    // the base's trailing 24h chains hold.
    static const bool sort_pairs = getenv("IGS_SORT_PAIRS") && atoi(getenv("IGS_SORT_PAIRS")) != 0;   // test knob (build_index)
    bool through_delta = !e->has_event_ids && !sort_pairs;
    if (through_delta) {
        {
            std::lock_guard<std::mutex> lk(e->view_mu);
            wait_for_readers(e, e->dv_spare, e->ingest_stream);
        }
        BuildSpec spec;
        spec.lo = e->base_n; spec.hi = e->n_raw; spec.n_accounts = n_accounts; spec.dst = &e->dv_spare;
        spec.base = view_of(e->ev, true); spec.n_before = before; spec.assume_time_ordered = true;
        // the most the delta holds before it is folded in (bounded: a huge threshold is not a reason to allocate for it)
        spec.reserve = std::min<uint64_t>(std::max<uint64_t>(e->delta_min, e->base_n / 8), 8u << 20) + n;
        uint32_t late[2];
        rc = build_index(e, spec, late);
        if (rc != IGS_OK) { e->n_raw = before; return rc; }
        if (late[0] != 0) {
            if (e->late_policy == IGS_LATE_REJECT) { e->n_raw = before; return reject_late(e, late[1] ? late[1] : late[0]); }
            through_delta = false;       // IGS_LATE_REORDER: file everything by timestamp in a new base
        } else {
            {
                std::lock_guard<std::mutex> lk(e->view_mu);
                std::swap(e->dv, e->dv_spare);
            }
            const uint64_t delta_n = e->n_raw - e->base_n;
            if (delta_n <= std::max<uint64_t>(e->delta_min, e->base_n / 8)) return IGS_OK;
            rc = rebuild_base(e, n_accounts, 0xFFFFFFFFull);          // compaction: nothing in it is new
            if (rc != IGS_OK) igs_set_error(e, "compaction of the event index failed; the appended events stay in the delta");
            return IGS_OK;               // the batch is in either way: base + delta still answer correctly
        }
    }
    rc = rebuild_base(e, n_accounts, before);
    if (rc != IGS_OK) e->n_raw = before;
    return rc;
}

// Everything an append allocates on demand, allocated now: room in the raw columns for the events until the next
// compaction, both sets of delta arrays, the spare set of base arrays for the compaction.
int igs_prepare_event_ingest(igs_engine* e, uint64_t max_batch) {
    if (!e->has_events) {
        igs_set_error(e, "igs_prepare_ingest needs a loaded history (igs_load_events, possibly with n = 0)");
        return IGS_ERR_STATE;
    }
    const uint32_t n_accounts = e->ev.n_accounts;
    // sized for the history growing by half: the compaction threshold grows with the base (base / 8), so arrays that
    // fit exactly today would be re-allocated right after the first compaction
    const uint64_t horizon = e->n_raw + e->n_raw / 2;
    const uint64_t threshold = std::min<uint64_t>(std::max<uint64_t>(e->delta_min, horizon / 8), 8u << 20);
    const uint64_t delta_cap = threshold + max_batch;
    if (horizon + delta_cap >= 0xFFFFFFFFull) {
        igs_set_error(e, "at most 2^32-2 events per shard");
        return IGS_ERR_INVALID;
    }
    {
        const int rc = reserve_raw(e, horizon + delta_cap, e->ingest_stream);
        if (rc != IGS_OK) return rc;
    }
    if (!e->has_event_ids) {
        int rc = ensure_index_capacity(e, e->dv_spare, delta_cap, n_accounts);
        if (rc == IGS_OK && e->dv.n == 0) rc = ensure_index_capacity(e, e->dv, delta_cap, n_accounts);
        if (rc != IGS_OK) return rc;
    }
    {
        const int rc = ensure_index_capacity(e, e->ev_spare, horizon + delta_cap, n_accounts);
        if (rc != IGS_OK) return rc;
    }
    // The published base becomes the spare set at the first compaction and would be re-allocated at the second:
    // move it into arrays of the same capacity now (a device-to-device copy; this is a set-up call).
    if (e->ev.cap_events < horizon + delta_cap + 1 && e->ev.n != 0) {
        igs_event_index big;
        int rc = ensure_index_capacity(e, big, horizon + delta_cap, n_accounts);
        if (rc != IGS_OK) { release_index_arrays(e, big); return rc; }
        cudaStream_t st = e->ingest_stream;
        const igs_event_index& x = e->ev;
        IGS_CUDA(e, cudaMemcpyAsync(big.seg, x.seg, ((size_t)x.n_accounts + 1) * 4, cudaMemcpyDeviceToDevice, st));
        IGS_CUDA(e, cudaMemcpyAsync(big.ts, x.ts, (size_t)x.n * 4, cudaMemcpyDeviceToDevice, st));
        IGS_CUDA(e, cudaMemcpyAsync(big.seq, x.seq, (size_t)x.n * 4, cudaMemcpyDeviceToDevice, st));
        IGS_CUDA(e, cudaMemcpyAsync(big.pre, x.pre, ((size_t)x.n + 1) * sizeof(igs_prefix_rec), cudaMemcpyDeviceToDevice, st));
        IGS_CUDA(e, cudaMemcpyAsync(big.tail, x.tail, (size_t)x.n * sizeof(igs_tail_rec), cudaMemcpyDeviceToDevice, st));
        IGS_CUDA(e, cudaStreamSynchronize(st));
        big.n = x.n;
        big.n_accounts = x.n_accounts;
        {
            std::lock_guard<std::mutex> lk(e->view_mu);
            std::swap(e->ev, big);                   // `big` now holds the old arrays and their readers
        }
        IGS_CUDA(e, cudaDeviceSynchronize());        // kernels still reading the old arrays
        std::lock_guard<std::mutex> lk(e->view_mu);
        release_index_arrays(e, big);
    }
    return IGS_OK;
}

int igs_build_event_index(igs_engine* e, const igs_event_cols* ev, const igs_event_ids* ids, uint64_t n64,
                          uint32_t n_accounts) {
    IGS_CUDA(e, cudaDeviceSynchronize());
    // The previous history's allocations (raw columns, index arrays, build scratch) are kept and reused when they
    // are large enough: allocating and freeing a dozen buffers of hundreds of MB costs tens of milliseconds, ten
    // times the kernels of the build.  They are released when the load fails and when the engine is destroyed.
    {
        std::lock_guard<std::mutex> lk(e->view_mu);
        e->has_events = false;
        e->ev.n = e->dv.n = 0;
    }
    const bool with_ids = ids != nullptr;
    if (with_ids != e->has_event_ids || (with_ids && !e->raw_device)) free_raw_events(e);   // different column set
    e->has_event_ids = with_ids;
    e->n_raw = e->base_n = 0;
    e->late_events = 0;
    int rc = append_raw(e, ev, ids, n64, e->ingest_stream);
    if (rc == IGS_OK) rc = rebuild_base(e, n_accounts, 0xFFFFFFFFull);
    if (rc != IGS_OK) igs_free_event_index(e);
    return rc;
}

int igs_launch_uniques(igs_engine* e, const igs_query* d_q, uint64_t n, igs_uniques* d_out, cudaStream_t stream) {
    if (!e->has_events || !e->has_event_ids) {
        igs_set_error(e, "no event history with identifiers loaded (igs_load_events_ids)");
        return IGS_ERR_STATE;
    }
    if (n == 0) return IGS_OK;
    const unsigned long long nb = (n + kThreads - 1) / kThreads;
    igs_history_reader rd(e);
    k_uniques<<<(unsigned)nb, kThreads, 0, stream>>>(rd.view, d_q, n, d_out);
    IGS_LAUNCHED(e);
    rd.done(stream);
    return IGS_OK;
}

int igs_launch_velocity(igs_engine* e, const igs_query* d_q, uint64_t n, igs_velocity* d_out,
                        cudaStream_t stream) {
    if (!e->has_events) {
        igs_set_error(e, "no event history loaded (igs_load_events)");
        return IGS_ERR_STATE;
    }
    if (n == 0) return IGS_OK;
    const unsigned long long nb = (n + kThreads - 1) / kThreads;
    igs_history_reader rd(e);
    k_velocity<<<(unsigned)nb, kThreads, 0, stream>>>(rd.view, d_q, n, d_out);
    IGS_LAUNCHED(e);
    rd.done(stream);
    return IGS_OK;
}

// Drops the events that can no longer influence any query at t >= now and rebuilds the index.  Stable
// compaction of the raw columns: the survivors keep their relative arrival order but are renumbered.
int igs_evict_event_history(igs_engine* e, uint32_t now, uint64_t* n_evicted) {
    if (n_evicted) *n_evicted = 0;
    if (!e->has_events) {
        igs_set_error(e, "no event history loaded (igs_load_events)");
        return IGS_ERR_STATE;
    }
    const uint32_t n = (uint32_t)e->n_raw;
    const uint32_t n_accounts = e->ev.n_accounts;
    if (n == 0) return IGS_OK;
    IGS_CUDA(e, cudaDeviceSynchronize());
    int rc = IGS_OK;
    if (e->base_n != e->n_raw) {                      // fold the delta in: the passes below walk one index
        rc = rebuild_base(e, n_accounts, 0xFFFFFFFFull);
        if (rc != IGS_OK) return rc;
    }
    cudaStream_t st = e->ingest_stream;
    const igs_event_index& x = e->ev;
    const uint32_t nb = (n + kThreads - 1) / kThreads;
    uint32_t *keep_from = nullptr, *keep = nullptr, *scr = nullptr, *last_keep = nullptr;
    void* fresh[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    auto bail = [&](cudaError_t err, const char* what) {
        igs_set_error(e, "%s failed: %s", what, cudaGetErrorString(err));
        rc = err == cudaErrorMemoryAllocation ? IGS_ERR_NOMEM : IGS_ERR_CUDA;
    };
    uint32_t kept = 0;
    do {
        cudaError_t err;
        if ((err = cudaMalloc(&keep_from, ((size_t)n_accounts + 1) * 4)) != cudaSuccess) { bail(err, "cudaMalloc"); break; }
        if ((err = cudaMalloc(&keep, (size_t)n * 4 + 4)) != cudaSuccess) { bail(err, "cudaMalloc"); break; }
        if ((err = cudaMalloc(&scr, scan_u32_scratch_elems(n) * 4)) != cudaSuccess) { bail(err, "cudaMalloc"); break; }
        last_keep = keep + n;
        cudaMemsetAsync(keep_from, 0, ((size_t)n_accounts + 1) * 4, st);
        k_ev_breaks<<<nb, kThreads, 0, st>>>(x.seg, x.ts, n, n_accounts, now, keep_from);
        k_ev_dead<<<(n_accounts + kThreads - 1) / kThreads, kThreads, 0, st>>>(x.seg, x.ts, n_accounts, now, keep_from);
        k_ev_keep<<<nb, kThreads, 0, st>>>(x.seg, x.seq, n, n_accounts, keep_from, keep);
        k_save_last_u32<<<1, 32, 0, st>>>(keep, n, last_keep);
        if ((err = cudaGetLastError()) != cudaSuccess) { bail(err, "eviction kernels"); break; }
        if ((rc = scan_u32(e, keep, n, scr, st)) != IGS_OK) break;
        uint32_t ex = 0, lk = 0;
        cudaMemcpyAsync(&ex, keep + (n - 1), 4, cudaMemcpyDeviceToHost, st);
        cudaMemcpyAsync(&lk, last_keep, 4, cudaMemcpyDeviceToHost, st);
        if ((err = cudaStreamSynchronize(st)) != cudaSuccess) { bail(err, "eviction count"); break; }
        kept = ex + lk;
        if (kept == n) break;                                     // nothing to drop
        const size_t cap = std::max<size_t>(kept, 1024);
        const size_t widths[6] = {4, 4, 8, 1, 8, 8};
        const int cols = e->has_event_ids ? 6 : 4;
        bool ok = true;
        for (int c = 0; c < cols && ok; ++c)
            if ((err = cudaMalloc(&fresh[c], cap * widths[c])) != cudaSuccess) { bail(err, "cudaMalloc"); ok = false; }
        if (!ok) break;
        k_ev_compact<<<nb, kThreads, 0, st>>>(e->raw_account, keep, last_keep, n, (uint32_t*)fresh[0]);
        k_ev_compact<<<nb, kThreads, 0, st>>>(e->raw_ts, keep, last_keep, n, (uint32_t*)fresh[1]);
        k_ev_compact<<<nb, kThreads, 0, st>>>(e->raw_amount, keep, last_keep, n, (int64_t*)fresh[2]);
        k_ev_compact<<<nb, kThreads, 0, st>>>(e->raw_type, keep, last_keep, n, (uint8_t*)fresh[3]);
        if (cols == 6) {
            k_ev_compact<<<nb, kThreads, 0, st>>>(e->raw_device, keep, last_keep, n, (uint64_t*)fresh[4]);
            k_ev_compact<<<nb, kThreads, 0, st>>>(e->raw_ip, keep, last_keep, n, (uint64_t*)fresh[5]);
        }
        if ((err = cudaGetLastError()) != cudaSuccess) { bail(err, "compaction"); break; }
        if ((err = cudaStreamSynchronize(st)) != cudaSuccess) { bail(err, "compaction"); break; }
        cudaFree(e->raw_account); cudaFree(e->raw_ts); cudaFree(e->raw_amount); cudaFree(e->raw_type);
        e->raw_account = (uint32_t*)fresh[0]; e->raw_ts = (uint32_t*)fresh[1];
        e->raw_amount = (int64_t*)fresh[2]; e->raw_type = (uint8_t*)fresh[3];
        if (cols == 6) {
            cudaFree(e->raw_device); cudaFree(e->raw_ip);
            e->raw_device = (uint64_t*)fresh[4]; e->raw_ip = (uint64_t*)fresh[5];
        }
        for (void*& f : fresh) f = nullptr;
        e->n_raw = kept;
        e->cap_raw = cap;
        rc = rebuild_base(e, n_accounts, 0xFFFFFFFFull);
    } while (false);
    for (void* f : fresh) if (f) cudaFree(f);
    cudaFree(keep_from); cudaFree(keep); cudaFree(scr);
    if (rc == IGS_OK && n_evicted) *n_evicted = n - kept;
    return rc;
}
