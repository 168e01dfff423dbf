// velocity.cuh — device-side lookup of one account's velocity block in the event index.
//
// The index (built by k_velocity.cu) holds the shard's events sorted by (account, ts, arrival)
// plus scan columns, so that "what would GetRealTimeFeatures return for this account at time t"
// (features/redis_store.go:60-116) is a handful of searches and prefix differences —
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
//
// Two levels.  The history is a BASE index plus a (possibly empty) DELTA index over the events appended
// since the base was built (igs_append_events): per account every delta event is ordered after every base
// event (appends with late events are refused or trigger a rebuild), the delta's chain columns continue the
// base's (carried in when the delta is built) and its distinct flags know the base's last (account, ts) group.
// A query whose prior events end in the base is answered from the base alone; one whose prior events reach
// into the delta adds the two levels' prefix differences and takes the chain columns from the delta.
//
// Searching.  A lookup is a chain of dependent loads, and on a hot account (10^5 .. 10^6 events) that chain is what
// it costs: five bisections of 20 probes each.  So (1) the position of the query among the account's events is a
// 4-ary search (three probes per round trip), with the "newer than everything" case — every live query — answered
// by the first round; (2) the four window bounds are searched together: four probes over the last eight events
// settle the typical account in ONE round trip, the rest is bisected in lock step — log2(length) round trips for
// all four bounds, not four times that.
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

// one level of the index
struct igs_index_view {
    const uint32_t* seg;             // [n_accounts+1]
    const uint32_t* ts;              // [n]
    const uint32_t* seq;             // [n] arrival index (global: base and delta share one numbering)
    const igs_prefix_rec* pre;       // [n+1]
    const igs_tail_rec* tail;        // [n]
    uint32_t n_accounts;
    uint32_t n;
};
// base + delta (delta.n == 0: no delta)
struct igs_history_view {
    igs_index_view base, delta;
};

// number of events in [s, e) ordered before (t, seq_limit), as a position: first one that is not.
// 4-ary search: three probes per round trip (independent loads) — half the dependent depth of a bisection, which
// is what a lookup in a long segment waits for.
__device__ __forceinline__ uint32_t igs_prior_end(const igs_index_view& ix, uint32_t s, uint32_t e, uint32_t t,
                                                  uint32_t seq_limit) {
    if (s == e) return s;
    auto before = [&](uint32_t tm, uint32_t sm) { return tm < t || (tm == t && sm < seq_limit); };
    uint32_t lo = s, hi = e;                      // answer in [lo, hi]
    {   // the common query is newer than everything the account has done: probe the last event together with the
        // first round of the search
        const uint32_t len = hi - lo;
        const uint32_t m1 = lo + len / 4, m2 = lo + len / 2, m3 = lo + (len / 4) * 3, ml = e - 1;
        const uint32_t t1 = __ldg(ix.ts + m1), t2 = __ldg(ix.ts + m2), t3 = __ldg(ix.ts + m3), tl = __ldg(ix.ts + ml);
        const uint32_t q1 = __ldg(ix.seq + m1), q2 = __ldg(ix.seq + m2), q3 = __ldg(ix.seq + m3), ql = __ldg(ix.seq + ml);
        if (before(tl, ql)) return e;
        hi = ml;
        if (!before(t1, q1)) hi = m1;
        else if (!before(t2, q2)) { lo = m1 + 1; hi = m2; }
        else if (!before(t3, q3)) { lo = m2 + 1; hi = m3; }
        else lo = m3 + 1;
    }
    while (lo < hi) {
        const uint32_t len = hi - lo;
        if (len <= 2) {
            const uint32_t tm = __ldg(ix.ts + lo), sm = __ldg(ix.seq + lo);
            if (!before(tm, sm)) return lo;
            ++lo;
            continue;
        }
        const uint32_t m1 = lo + len / 4, m2 = lo + len / 2, m3 = lo + (len / 4) * 3;
        const uint32_t t1 = __ldg(ix.ts + m1), t2 = __ldg(ix.ts + m2), t3 = __ldg(ix.ts + m3);
        const uint32_t q1 = __ldg(ix.seq + m1), q2 = __ldg(ix.seq + m2), q3 = __ldg(ix.seq + m3);
        if (!before(t1, q1)) hi = m1;
        else if (!before(t2, q2)) { lo = m1 + 1; hi = m2; }
        else if (!before(t3, q3)) { lo = m2 + 1; hi = m3; }
        else lo = m3 + 1;
    }
    return lo;
}

// window bounds of the prior events [s, u) of one level
struct igs_bounds {
    uint32_t lo60, lo300, lo3600, lo86400;
    uint32_t alive;   // first position that survived the last 1h trim (ts > L-3600); s when it cannot bind
};
// The four lower bounds (first position with ts >= now - W) are searched TOGETHER:
//   1. four probes at u-1, u-2, u-4, u-8 — independent loads in one or two sectors of ts[] — bracket every bound
//      that lies within the last eight events (the typical account: one round trip for all four bounds);
//   2. what is left is bisected in lock step, one probe per bound and round trip, so a long segment costs
//      log2(length) dependent loads for the four bounds together instead of four times that.
// `last_is_now`: the newest prior event (of the whole history, both levels) has ts == now.  Only then can the
// trim bound L-3599 exceed the 1h window bound now-3600 (quirk Q4); otherwise alive <= lo3600 and never binds.
__device__ __forceinline__ igs_bounds igs_window_bounds(const uint32_t* __restrict__ ts, uint32_t s, uint32_t u,
                                                        long long now, bool last_is_now) {
    igs_bounds b;
    b.alive = s;
    if (u == s) { b.lo60 = b.lo300 = b.lo3600 = b.lo86400 = s; return b; }
    const long long w[4] = {60, 300, 3600, 86400};
    uint32_t key[4], left[4], right[4];
    const uint32_t len = u - s;
    const uint32_t p1 = u - 1, p2 = len > 2 ? u - 2 : s, p4 = len > 4 ? u - 4 : s, p8 = len > 8 ? u - 8 : s;
    const uint32_t t1 = __ldg(ts + p1), t2 = __ldg(ts + p2), t4 = __ldg(ts + p4), t8 = __ldg(ts + p8);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const long long k = now - w[j];
        key[j] = k > 0 ? (uint32_t)k : 0u;        // now < 2^32: the key fits; a key <= 0 selects everything
        if (t1 < key[j]) { left[j] = right[j] = u; }
        else if (t8 >= key[j]) { left[j] = s; right[j] = p8; }
        else if (t4 >= key[j]) { left[j] = p8 + 1; right[j] = p4; }
        else if (t2 >= key[j]) { left[j] = p4 + 1; right[j] = p2; }
        else { left[j] = p2 + 1; right[j] = p1; }
    }
    // the bounds are nested (a longer window starts no later): a settled bound caps the ones after it
#pragma unroll
    for (int j = 1; j < 4; ++j)
        if (right[j] > right[j - 1]) { right[j] = right[j - 1]; if (left[j] > right[j]) left[j] = right[j]; }
    // bisection in lock step: one probe per bound and round trip.  (Three probes per bound — a 4-ary search, ten
    // round trips instead of twenty on C3's longest segment — was measured: 12 loads per round cost more in traffic
    // and registers than the shorter chain saves, 0.363 -> 0.396 ms per 1M queries, fused path 1.67 -> 1.74 ms.)
    while ((left[0] < right[0]) | (left[1] < right[1]) | (left[2] < right[2]) | (left[3] < right[3])) {
        uint32_t mid[4], tm[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            mid[j] = left[j] + ((right[j] - left[j]) >> 1);
            tm[j] = left[j] < right[j] ? __ldg(ts + mid[j]) : 0u;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (left[j] < right[j]) {
                if (tm[j] < key[j]) left[j] = mid[j] + 1; else right[j] = mid[j];
            }
    }
    b.lo60 = left[0]; b.lo300 = left[1]; b.lo3600 = left[2]; b.lo86400 = left[3];
    if (last_is_now) {                            // now-3599 lies between the 1h and the 5m bound
        uint32_t lo = b.lo3600, hi = b.lo300;
        const uint32_t k = now - 3599 > 0 ? (uint32_t)(now - 3599) : 0u;
        while (lo < hi) {
            const uint32_t mid = lo + ((hi - lo) >> 1);
            if (__ldg(ts + mid) < k) lo = mid + 1; else hi = mid;
        }
        b.alive = lo;
    }
    return b;
}

// adds one level's contribution: prefix differences between position u and the bounds
__device__ __forceinline__ void igs_velocity_accumulate(const igs_prefix_rec* __restrict__ pre, uint32_t u,
                                                        const igs_bounds& b, igs_velocity* v) {
    const igs_prefix_rec pu = pre[u];
    const igs_prefix_rec p60 = pre[b.lo60];
    const igs_prefix_rec p3600 = pre[b.lo3600];
    const igs_prefix_rec p86400 = pre[b.lo86400];
    const uint32_t a60 = b.lo60 > b.alive ? b.lo60 : b.alive;
    const uint32_t a300 = b.lo300 > b.alive ? b.lo300 : b.alive;
    const uint32_t a3600 = b.lo3600 > b.alive ? b.lo3600 : b.alive;
    const uint32_t d60 = a60 == b.lo60 ? p60.dcnt : pre[a60].dcnt;
    const uint32_t d300 = a300 == b.lo60 ? p60.dcnt : (a300 == b.lo3600 ? p3600.dcnt : pre[a300].dcnt);
    const uint32_t d3600 = a3600 == b.lo3600 ? p3600.dcnt : pre[a3600].dcnt;
    v->tx_count_1m += pu.dcnt - d60;
    v->tx_count_5m += pu.dcnt - d300;
    v->tx_count_1h += pu.dcnt - d3600;
#pragma unroll
    for (int ty = 0; ty < 2; ++ty) {
        v->win[ty][0].count += pu.cnt[ty] - p60.cnt[ty];
        v->win[ty][0].sum += pu.sum[ty] - p60.sum[ty];
        v->win[ty][1].count += pu.cnt[ty] - p3600.cnt[ty];
        v->win[ty][1].sum += pu.sum[ty] - p3600.sum[ty];
        v->win[ty][2].count += pu.cnt[ty] - p86400.cnt[ty];
        v->win[ty][2].sum += pu.sum[ty] - p86400.sum[ty];
    }
}

// the chain columns of the newest prior event (position `last` of level ix)
__device__ __forceinline__ void igs_velocity_tail(const igs_index_view& ix, uint32_t last, long long now, long long L,
                                                  igs_velocity* v, igs_uniques* uq) {
    const long long gap = now - L;
    union __align__(32) { igs_tail_rec r; uint4 q[2]; } tlu;
    {
        const uint4* src = reinterpret_cast<const uint4*>(ix.tail + last);
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

__device__ __forceinline__ void igs_velocity_lookup(const igs_history_view& hv, uint32_t account, uint32_t t,
                                                    uint32_t seq_limit, igs_velocity* v, igs_uniques* uq = nullptr) {
    if (uq) uq->devices_24h = uq->ips_24h = 0;
    v->tx_sum_1h = 0;
    v->tx_count_1m = v->tx_count_5m = v->tx_count_1h = 0;
    v->time_since_last_tx = v->session_duration = v->n_prior = 0;
#pragma unroll
    for (int ty = 0; ty < 2; ++ty)
#pragma unroll
        for (int w = 0; w < 3; ++w) { v->win[ty][w].sum = 0; v->win[ty][w].count = 0; v->win[ty][w].pad = 0; }
    const igs_index_view& bx = hv.base;
    const igs_index_view& dx = hv.delta;
    uint32_t sb = 0, eb = 0, sd = 0, ed = 0;
    if (account < bx.n_accounts) { sb = __ldg(bx.seg + account); eb = __ldg(bx.seg + account + 1); }
    if (dx.n != 0 && account < dx.n_accounts) { sd = __ldg(dx.seg + account); ed = __ldg(dx.seg + account + 1); }
    const long long now = t;
    const uint32_t ud = igs_prior_end(dx, sd, ed, t, seq_limit);
    if (ud == sd) {                                     // the prior events end in the base
        const uint32_t ub = igs_prior_end(bx, sb, eb, t, seq_limit);
        if (ub == sb) return;
        v->n_prior = ub - sb;
        const long long L = __ldg(bx.ts + ub - 1);
        const igs_bounds b = igs_window_bounds(bx.ts, sb, ub, now, L == now);
        igs_velocity_accumulate(bx.pre, ub, b, v);
        igs_velocity_tail(bx, ub - 1, now, L, v, uq);
        return;
    }
    // the prior events reach into the delta: every base event of the account is prior too
    v->n_prior = (eb - sb) + (ud - sd);
    const long long L = __ldg(dx.ts + ud - 1);
    const igs_bounds d = igs_window_bounds(dx.ts, sd, ud, now, L == now);
    igs_velocity_accumulate(dx.pre, ud, d, v);
    if (eb > sb && d.lo86400 == sd) {                   // some window reaches back into the base
        igs_bounds b = igs_window_bounds(bx.ts, sb, eb, now, L == now);
        // a bound that fell inside the delta excludes the base altogether for that window
        if (d.lo60 != sd) b.lo60 = eb;
        if (d.lo300 != sd) b.lo300 = eb;
        if (d.lo3600 != sd) b.lo3600 = eb;
        if (d.alive != sd) b.alive = eb;
        igs_velocity_accumulate(bx.pre, eb, b, v);
    }
    igs_velocity_tail(dx, ud - 1, now, L, v, uq);       // chain columns were carried over from the base at build time
}

// views of the engine's loaded history (n_accounts = 0 when nothing is loaded), taken under the engine's view lock
struct igs_history_view igs_history_view_of(igs_engine* e);
