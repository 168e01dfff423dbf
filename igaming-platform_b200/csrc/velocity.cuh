// velocity.cuh — device-side lookup of one account's velocity block in the event index.
//
// The index (built by k_velocity.cu) holds the shard's events sorted by (account, ts, arrival)
// plus scan columns, so that "what would GetRealTimeFeatures return for this account at time t"
// (features/redis_store.go:60-116) is a handful of binary searches and prefix differences —
// O(log events-of-the-account) per query, independent of the history size.
//
// Closed form (proven equal to a command-level Redis emulation by the CPU tests):
//   prior events  = events of the account ordered before (t, seq_limit); L = ts of the last one
//   ZSET members  = distinct (ts, amount) with ts > L-3600           (ZADD dedupe :127-130, trim :132)
//   cW            = members with ts >= t-W                           (ZCOUNT, inclusive :72-74)
//   tx_sum:1h     = sum over the trailing chain with gaps <= 3600 iff t-L <= 3600   (:136-138)
//   last_tx       = L iff t-L <= 7d (:155);  session_start = head of the gap<=1800 chain iff
//                   t-L <= 1800 (:158-160)
//   extension     = per-type (bet, deposit) count / sum over [t-W, t], W = 1m, 1h, 24h, multiset
//   uniques       = distinct identifiers of the trailing chain (gaps <= 86400) of the prior events that
//                   carried one, iff t - (ts of the last of them) <= 86400   (PFADD/EXPIRE :141-152)
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "igs.h"

struct igs_engine;

// exclusive prefix columns at sorted position i (i = 0..n), one 32-byte sector per lookup
struct __align__(32) igs_prefix_rec {
    uint32_t dcnt;      // distinct (account, ts, amount) first occurrences before i
    uint32_t cnt[2];    // events of type bet / deposit before i
    uint32_t pad;
    int64_t  sum[2];    // their amounts
};
// inclusive chain columns of sorted event i — one 32-byte sector, so the identifier columns ride
// along with the load the chain columns need anyway
struct __align__(32) igs_tail_rec {
    int64_t  chain_sum;   // sum over the maximal chain (consecutive gaps <= 3600) ending at i
    uint32_t sess_start;  // ts of the first event of the gap<=1800 chain ending at i
    // PFADD + EXPIRE 24h keys (features/redis_store.go:141-152) as of event i, per key: ts of the
    // account's last event <= i that carried an identifier, and the number of distinct identifiers in
    // its chain (gaps <= 86400 between consecutive such events).  cnt == 0: no such event.
    uint32_t dev_ts, ip_ts;
    uint16_t dev_cnt, ip_cnt;
    uint32_t pad[2];
};
static_assert(sizeof(igs_tail_rec) == 32, "one sector per tail record");

// a velocity block that can be moved with 16-byte loads/stores
union __align__(16) igs_velocity_a16 {
    igs_velocity v;
    uint4 q[sizeof(igs_velocity) / 16];
    __device__ igs_velocity_a16() {}
};

struct igs_index_view {
    const uint32_t* seg;             // [n_accounts+1]
    const uint32_t* ts;              // [n]
    const uint32_t* seq;             // [n]
    const igs_prefix_rec* pre;       // [n+1]
    const igs_tail_rec* tail;        // [n]
    uint32_t n_accounts;
    uint32_t n;
};

// first i in [lo, hi) with ts[i] >= key  (key may be negative: then lo)
__device__ __forceinline__ uint32_t igs_lower_ts(const uint32_t* __restrict__ ts, uint32_t lo,
                                                 uint32_t hi, long long key) {
    if (key <= 0) return lo;
    if (key > 0xFFFFFFFFll) return hi;
    const uint32_t k = (uint32_t)key;
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if (__ldg(ts + mid) < k) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ void igs_velocity_lookup(const igs_index_view& ix, uint32_t account,
                                                    uint32_t t, uint32_t seq_limit, igs_velocity* v,
                                                    igs_uniques* uq = nullptr) {
    if (uq) uq->devices_24h = uq->ips_24h = 0;
    v->tx_sum_1h = 0;
    v->tx_count_1m = v->tx_count_5m = v->tx_count_1h = 0;
    v->time_since_last_tx = v->session_duration = v->n_prior = 0;
#pragma unroll
    for (int ty = 0; ty < 2; ++ty)
#pragma unroll
        for (int w = 0; w < 3; ++w) { v->win[ty][w].sum = 0; v->win[ty][w].count = 0; v->win[ty][w].pad = 0; }
    if (account >= ix.n_accounts) return;
    const uint32_t s = __ldg(ix.seg + account), e = __ldg(ix.seg + account + 1);
    // u = number of prior events: first position whose (ts, seq) is not < (t, seq_limit)
    uint32_t lo = s, hi = e;
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        const uint32_t tm = __ldg(ix.ts + mid);
        const bool before = tm < t || (tm == t && __ldg(ix.seq + mid) < seq_limit);
        if (before) lo = mid + 1; else hi = mid;
    }
    const uint32_t u = lo;
    if (u == s) return;
    v->n_prior = u - s;
    const long long now = t;
    const long long L = __ldg(ix.ts + u - 1);
    const uint32_t lo60 = igs_lower_ts(ix.ts, s, u, now - 60);
    const uint32_t lo300 = igs_lower_ts(ix.ts, s, lo60, now - 300);
    const uint32_t lo3600 = igs_lower_ts(ix.ts, s, lo300, now - 3600);
    const uint32_t lo86400 = igs_lower_ts(ix.ts, s, lo3600, now - 86400);
    const uint32_t alive = igs_lower_ts(ix.ts, s, u, L - 3599);   // ts > L-3600
    const igs_prefix_rec pu = ix.pre[u];
    const igs_prefix_rec p60 = ix.pre[lo60];
    const igs_prefix_rec p3600 = ix.pre[lo3600];
    const igs_prefix_rec p86400 = ix.pre[lo86400];
    const uint32_t a300 = lo300 > alive ? lo300 : alive;
    const uint32_t a3600 = lo3600 > alive ? lo3600 : alive;
    v->tx_count_1m = pu.dcnt - (lo60 >= alive ? p60.dcnt : ix.pre[alive].dcnt);
    v->tx_count_5m = pu.dcnt - ix.pre[a300].dcnt;
    v->tx_count_1h = pu.dcnt - (a3600 == lo3600 ? p3600.dcnt : ix.pre[a3600].dcnt);
#pragma unroll
    for (int ty = 0; ty < 2; ++ty) {
        v->win[ty][0].count = pu.cnt[ty] - p60.cnt[ty];
        v->win[ty][0].sum = pu.sum[ty] - p60.sum[ty];
        v->win[ty][1].count = pu.cnt[ty] - p3600.cnt[ty];
        v->win[ty][1].sum = pu.sum[ty] - p3600.sum[ty];
        v->win[ty][2].count = pu.cnt[ty] - p86400.cnt[ty];
        v->win[ty][2].sum = pu.sum[ty] - p86400.sum[ty];
    }
    const long long gap = now - L;
    union __align__(16) { igs_tail_rec r; uint4 q[2]; } tlu;
    {
        const uint4* src = reinterpret_cast<const uint4*>(ix.tail + (u - 1));
        tlu.q[0] = __ldg(src);
        if (uq) tlu.q[1] = __ldg(src + 1);
    }
    const igs_tail_rec& tl = tlu.r;
    if (uq) {
        if (tl.dev_cnt && now - (long long)tl.dev_ts <= 86400) uq->devices_24h = tl.dev_cnt;
        if (tl.ip_cnt && now - (long long)tl.ip_ts <= 86400) uq->ips_24h = tl.ip_cnt;
    }
    if (gap <= 3600) v->tx_sum_1h = tl.chain_sum;
    if (gap <= 7 * 24 * 3600) v->time_since_last_tx = (uint32_t)gap;
    if (gap <= 1800) v->session_duration = (uint32_t)(now - (long long)tl.sess_start);
}

// view of the engine's loaded index (n_accounts = 0 when no history is loaded)
struct igs_index_view igs_index_view_of(const igs_engine* e);
