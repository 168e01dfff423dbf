// k_score.cu — fused rules + tree-ensemble + blend kernel (SURVEY K5+K6+K7, and K3-lookup+K4 for the
// fused path), sm_100a.
//
// One persistent CTA per SM, W warps, one transaction per lane (32 W tx per tile).
//   * Each warp stages its 32 x 64 fp32 feature rows into a private smem tile laid out
//     [feature][lane]: lane l only ever touches bank l, so the data-dependent x[feat] gather of
//     the tree walk is bank-conflict free whatever features the 32 lanes ask for.
//   * The ensemble streams through smem in chunks of C trees moved by one cp.async.bulk (TMA bulk
//     copy, SASS UBLKCP) each, double buffered behind mbarriers, so the L2->smem traffic overlaps
//     the walk of the previous chunk.
//   * Tree levels are packed in pairs (igs_internal.cuh): 5 shared-memory wavefronts per two levels
//     instead of 6, and levels 0..5 are conflict free.
//   * A lane walks 5 trees at once (independent LDS->LDS->FSET chains) to cover the shared-memory
//     latency; the index arithmetic is written for the FMA pipe (IMAD) because the ALU pipe and
//     the shared-memory pipe are the two limiters (profiles/).
//   * Leaves are added to the margin in tree-index order in fp32 — the oracle's order — so the
//     margin, the logistic (igs_sigmoid), the fp64 blend and the integer score are bit-identical
//     to the CPU restatement.
//   * Launch shapes (warps x chunk trees x ILP [x CTAs per SM]): 12 x 20 x 5 for pre-assembled rows; the fused
//     path runs two independent 8 x 10 x 5 CTAs per SM so that one CTA's lookup phase overlaps the other's walk.
// Two input sources share the kernel body:
//   kSrcFeatures (path A): pre-assembled normalised 64-float rows + the 48-byte rule sideband.
//   kSrcTx       (path B): 32-byte tx records; the lane looks its account up in the event index
//                (velocity.cuh), gathers the player row, runs extractFeatures + the scoring->ml
//                mapping + Normalize in registers and writes the 64 floats straight into the
//                smem tile — the assembled row never touches HBM (SURVEY K3-lookup + K4 fused).
// Replaces: ScoringEngine.Score (scoring/engine.go:262-323) incl. extractFeatures (:326-417),
// ONNXModel.PredictBatch (ml/onnx_model.go:311-326), Normalize (:169-184).
#include "idset.cuh"
#include "igs_internal.cuh"
#include "igs_math.cuh"
#include "onesweep.cuh"
#include "sortscan.cuh"
#include "velocity.cuh"

#include <stdlib.h>
#include <algorithm>
#include <string.h>

namespace {

constexpr int kFeatTileFloats = IGS_NUM_FEATURES * 32;           // per warp
constexpr int kFeatTileBytes = kFeatTileFloats * 4;              // 8 KB

// Launch shape: W warps per CTA (one tx per lane -> 32 W tx per tile), tree chunks of C trees.
// smem = W x 8 KB feature tiles + 2 x C x 3 KB chunk buffers.
template <int W, int C, int I = 5, int B = 1>
struct Shape {
    static constexpr int kIlp = I;        // trees walked at once per lane
    static constexpr int kCtasPerSm = B;  // independent CTAs per SM (their phases interleave)
    static constexpr int kWarps = W;
    static constexpr int kThreads = W * 32;
    static constexpr int kTileTx = W * 32;
    static constexpr int kChunkTrees = C;
    static constexpr int kChunkBytes = C * IGS_DEV_TREE_BYTES;
    static constexpr int kSmemBytes = W * kFeatTileBytes + 2 * kChunkBytes + 64;
    static_assert(C % I == 0, "chunk must be a whole number of ILP groups");
    static_assert(B * (kSmemBytes + 1024) <= 227 * 1024, "smem budget");
};

constexpr int kSrcFeatures = 0;
constexpr int kSrcTx = 1;

struct ScoreParams {
    igs_config cfg;
    // path A
    const float* features;
    const igs_rule_row* rows;
    // path B
    const igs_tx* tx;
    const igs_player_row* players;
    unsigned long long n_players;
    igs_history_view ix;
    float* features_out;
    igs_velocity* velocity_out;
    const uint64_t* perm;          // optional: low 32 bits = slot -> row, the batch visited in account order (see launch)
    const uint32_t* tx_ready;      // optional (k_score_ws only): tx[] is still being copied in, see igs_tx_arrival
    uint32_t tx_chunk_rows;
    const igs_tx_ids* tx_ids;      // optional: rule 8 evaluated against the blacklist sets
    igs_idset_view bl[3];
    int uniques_from_index;        // the history carries identifiers: ignore the tx's unique counts
    // common
    igs_result* out;
    uint8_t* leaf_idx;
    unsigned long long n;
    const uint8_t* model;
    const uint8_t* model_ws;       // the same ensemble with the feature indices >= 42 mapped onto row 42 (k_score_ws)
    unsigned n_trees;
    float base_score;
    unsigned post;
};

// What applyRules reads, in registers.
struct RuleIn {
    long long amount, total_deposits, total_withdrawals;
    unsigned tx_count_1m, tslt, deposit_count, uniq_dev, uniq_ip, bits;
    int account_age, tx_type;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
        "l"(src), "r"(bytes), "r"(bar)
        : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            " selp.u32 %0, 1, 0, p;\n}\n"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!ok);
}

// --- instruction-mix helpers -------------------------------------------------------------------
// The walk is bound by the ALU pipe and the shared-memory pipe together, so the index arithmetic is
// written to land on the FMA pipe (IMAD) and the compare produces a mask (FSET) instead of a
// predicate + select.
__device__ __forceinline__ uint32_t fset_gtu(float x, float thr) {     // 0xFFFFFFFF iff !(x <= thr): BRANCH_LEQ, NaN -> right
    uint32_t m;
    asm("set.gtu.u32.f32 %0, %1, %2;" : "=r"(m) : "f"(x), "f"(thr));
    return m;
}
__device__ __forceinline__ uint32_t imad(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
template <int kOff>
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2+%3];" : "=r"(v.x), "=r"(v.y) : "r"(addr), "n"(kOff));
    return v;
}
template <int kOff>
__device__ __forceinline__ float lds32f(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v) : "r"(addr), "n"(kOff));
    return v;
}

// One level pair for one tree: even level `kD` (8-byte node) and odd level kD+1 (4-byte threshold).
// n = minus the node index inside the level (so that "2*index + went_right" is one IMAD: n' = 2n + mask).
template <int kD>
__device__ __forceinline__ void walk_pair(uint32_t tree_addr, uint32_t tile_lane_addr, uint32_t& n) {
    const uint2 nd = lds64<igs_dev_level_off(kD)>(imad(n, 0xFFFFFFF8u /* -8 */, tree_addr));
    const float x0 = lds32f<0>((nd.x & 0x1F80u) + tile_lane_addr);
    const uint32_t m0 = fset_gtu(x0, __uint_as_float(nd.y));
    const uint32_t sh = imad(m0, 0xFFFFFFFAu /* -6 */, 6u);             // 6: left child's field, 12: right child's
    n = imad(n, 2u, m0);
    const uint32_t child = ((nd.x >> sh) & 0x1F80u) + tile_lane_addr;
    const float thr1 = lds32f<igs_dev_level_off(kD + 1)>(imad(n, 0xFFFFFFFCu /* -4 */, tree_addr));
    const float x1 = lds32f<0>(child);
    n = imad(n, 2u, fset_gtu(x1, thr1));
}

// Walk `trees` trees of the chunk at smem address `cb` for this lane's transaction.  `tile_lane_addr` =
// smem address of this warp's [64][32] fp32 feature tile offset by lane.  Adds leaves to *margin in
// tree order.  Levels are visited in pairs (igs_internal.cuh): LDS.64 {feats, thr} at the even level,
// LDS.32 thr at the odd level whose feature came with the parent.
template <bool kLeaves, int kIlp>
__device__ __forceinline__ void walk_chunk(uint32_t cb, int trees, uint32_t tile_lane_addr, float* margin,
                                           uint8_t* leaf_out) {
    float m = *margin;
    int t = 0;
    for (; t + kIlp <= trees; t += kIlp) {
        const uint32_t tb = cb + t * IGS_DEV_TREE_BYTES;
        uint32_t n[kIlp];
#pragma unroll
        for (int k = 0; k < kIlp; ++k) n[k] = 0;
#pragma unroll
        for (int k = 0; k < kIlp; ++k) walk_pair<0>(tb + k * IGS_DEV_TREE_BYTES, tile_lane_addr, n[k]);
#pragma unroll
        for (int k = 0; k < kIlp; ++k) walk_pair<2>(tb + k * IGS_DEV_TREE_BYTES, tile_lane_addr, n[k]);
#pragma unroll
        for (int k = 0; k < kIlp; ++k) walk_pair<4>(tb + k * IGS_DEV_TREE_BYTES, tile_lane_addr, n[k]);
#pragma unroll
        for (int k = 0; k < kIlp; ++k) walk_pair<6>(tb + k * IGS_DEV_TREE_BYTES, tile_lane_addr, n[k]);
#pragma unroll
        for (int k = 0; k < kIlp; ++k) {
            m = __fadd_rn(m, lds32f<IGS_DEV_LEAF_OFF>(imad(n[k], 0xFFFFFFFCu, tb + k * IGS_DEV_TREE_BYTES)));
            if (kLeaves) leaf_out[t + k] = (uint8_t)(0u - n[k]);
        }
    }
    for (; t < trees; ++t) {
        const uint32_t tb = cb + t * IGS_DEV_TREE_BYTES;
        uint32_t n = 0;
        walk_pair<0>(tb, tile_lane_addr, n);
        walk_pair<2>(tb, tile_lane_addr, n);
        walk_pair<4>(tb, tile_lane_addr, n);
        walk_pair<6>(tb, tile_lane_addr, n);
        m = __fadd_rn(m, lds32f<IGS_DEV_LEAF_OFF>(imad(n, 0xFFFFFFFCu, tb)));
        if (kLeaves) leaf_out[t] = (uint8_t)(0u - n);
    }
    *margin = m;
}

// mockPredict, ml/onnx_model.go:258-308, on the already-normalised tile column of this lane.
template <bool kCompact = false>
__device__ __forceinline__ double mock_predict(const float* tile_lane) {
    auto v = [&](int i) { return tile_lane[(kCompact ? igs_ws_row(i) : i) * 32]; };
    double s = 0.0;
    if (v(0) > 0.5f) s = __dadd_rn(s, 0.2);                        // :262
    if (v(2) > 0.5f) s = __dadd_rn(s, 0.15);                       // :265
    if (v(5) > 0.3f) s = __dadd_rn(s, 0.15);                       // :270
    if (v(6) > 0.25f) s = __dadd_rn(s, 0.1);                       // :273
    if (v(19) > 0.f || v(20) > 0.f) s = __dadd_rn(s, 0.15);        // :278
    if (v(21) > 0.f) s = __dadd_rn(s, 0.25);                       // :281
    if (v(9) < 0.02f && v(26) > 0.5f) s = __dadd_rn(s, 0.2);       // :286
    if (v(25) > 0.f) s = __dadd_rn(s, 0.15);                       // :291
    if (v(15) < 0.01f && v(28) > 0.f) {                            // :296
        if (v(11) > __fmul_rn(v(10), 0.8f)) s = __dadd_rn(s, 0.2); // :297 (f32 multiply)
    }
    if (s > 1.0) s = 1.0;                                          // :303
    return s;
}

// Path B staging for one lane: velocity lookup + player gather + extractFeatures
// (scoring/engine.go:326-417) + scoring->ml mapping (ToSlice order, ml/onnx_model.go:133-166) +
// Normalize (:169-184) + extension block.  Writes the lane's column of the [64][32] tile.
// kCompact: the tile of the warp-specialised kernel — IGS_WS_ROWS rows, the features that are constant zero on this
// path share the last one (igs_ws_row); the engine's second model image uses the same row numbers.
template <bool kCompact = false>
__device__ __forceinline__ void assemble_tx(const ScoreParams& p, unsigned long long row, bool in_range,
                                            float* tile_lane /* my_tile + lane */, RuleIn* r) {
    igs_tx t;
    {
        uint4* dst = reinterpret_cast<uint4*>(&t);
        const uint4* src = reinterpret_cast<const uint4*>(p.tx + row);
        const uint4 z = make_uint4(0, 0, 0, 0);
        dst[0] = in_range ? __ldg(src + 0) : z;
        dst[1] = in_range ? __ldg(src + 1) : z;
    }
    // Blacklist.Check (scoring/engine.go:469-475): the probes are in flight while the velocity lookup runs
    igs_idset_probe3 blp;
    const bool check_bl = p.tx_ids != nullptr && in_range;
    if (check_bl) {
        const igs_tx_ids id = p.tx_ids[row];
        igs_idset_probe_begin(p.bl, id.device, id.fingerprint, id.ip, &blp);
    }
    igs_velocity_a16 vb;
    igs_uniques uq;
    igs_velocity_lookup(p.ix, in_range ? t.account : 0xFFFFFFFFu, t.ts, t.seq_limit, &vb.v, &uq);
    const igs_velocity& v = vb.v;
    // PFCOUNT :80-81 — from the history when it carries identifiers, else the host's counts
    const unsigned uniq_dev = p.uniques_from_index ? uq.devices_24h : t.unique_devices_24h;
    const unsigned uniq_ip = p.uniques_from_index ? uq.ips_24h : t.unique_ips_24h;
    unsigned tbits = t.bits;
    if (check_bl && igs_idset_probe_end(p.bl, blp)) tbits |= IGS_BIT_BLACKLISTED;
    if (p.velocity_out && in_range) {
        uint4* dst = reinterpret_cast<uint4*>(p.velocity_out + row);
#pragma unroll
        for (int k = 0; k < (int)(sizeof(igs_velocity) / 16); ++k) dst[k] = vb.q[k];
    }
    // BatchFeatures row (GetBatchFeatures, scoring/engine.go:362); absent -> zero features (:363-366).  (Requesting it
    // before the lookup would hide one round trip but keeps 24 registers alive across it: measured slower.)
    const bool have_player = in_range && (unsigned long long)t.account < p.n_players;
    union __align__(16) { igs_player_row r; uint4 q[sizeof(igs_player_row) / 16]; } pl;
    {
        const uint4* src = reinterpret_cast<const uint4*>(p.players + (have_player ? t.account : 0));
        const uint4 z = make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int k = 0; k < (int)(sizeof(igs_player_row) / 16); ++k) pl.q[k] = have_player ? __ldg(src + k) : z;
    }
    const long long now = t.ts;
    int account_age = 0;
    float win_rate = 0.f;
    bool bonus_only = false;
    if (have_player) {
        account_age = (int)((now - pl.r.account_created_at) / 86400);          // :374
        if (pl.r.bet_count > 0)                                                // :379-381
            win_rate = __double2float_rn(__ddiv_rn((double)pl.r.win_count, (double)pl.r.bet_count));
        bonus_only = pl.r.bonus_claim_count > 3 && pl.r.total_deposits < 5000;  // :384-386
    }
    float avg1h = 0.f;
    if (v.tx_count_1h > 0)                                                     // :412-414
        avg1h = __double2float_rn(__ddiv_rn((double)v.tx_sum_1h, (double)v.tx_count_1h));
    const long long net = pl.r.total_deposits - pl.r.total_withdrawals;        // :370
    auto put = [&](int i, float x) {
        if (!kCompact) tile_lane[i * 32] = x;
        else if (igs_ws_row(i) != IGS_WS_ROWS - 1) tile_lane[igs_ws_row(i) * 32] = x;     // constant zeros: written once below
    };
    if (kCompact) tile_lane[(IGS_WS_ROWS - 1) * 32] = 0.f;
    put(0, igs_min_max(__uint2float_rn(v.tx_count_1m), 20.f));
    put(1, igs_min_max(__uint2float_rn(v.tx_count_5m), 50.f));
    put(2, igs_min_max(__uint2float_rn(v.tx_count_1h), 200.f));
    put(3, igs_log_transform(__ll2float_rn(v.tx_sum_1h)));
    put(4, avg1h);
    put(5, igs_min_max(__uint2float_rn(uniq_dev), 10.f));
    put(6, igs_min_max(__uint2float_rn(uniq_ip), 20.f));
    put(7, 0.f);                                   // IPCountryChanges: never set by the reference
    put(8, 0.f);                                   // DeviceAge: never set
    put(9, igs_min_max(__int2float_rn(account_age), 365.f));
    put(10, igs_log_transform(__ll2float_rn(pl.r.total_deposits)));
    put(11, igs_log_transform(__ll2float_rn(pl.r.total_withdrawals)));
    put(12, __ll2float_rn(net));
    put(13, __int2float_rn(pl.r.deposit_count));
    put(14, __int2float_rn(pl.r.withdraw_count));
    put(15, igs_min_max(__uint2float_rn(v.time_since_last_tx), 86400.f));
    put(16, __uint2float_rn(v.session_duration));
    put(17, __double2float_rn(pl.r.avg_bet_size));
    put(18, win_rate);
    put(19, (t.bits & IGS_BIT_VPN) ? 1.f : 0.f);
    put(20, (t.bits & IGS_BIT_PROXY) ? 1.f : 0.f);
    put(21, (t.bits & IGS_BIT_TOR) ? 1.f : 0.f);
    put(22, 0.f);                                  // DisposableEmail: never set
    put(23, __int2float_rn(pl.r.bonus_claim_count));
    put(24, __double2float_rn(pl.r.bonus_wager_complete));
    put(25, bonus_only ? 1.f : 0.f);
    put(26, igs_log_transform(__ll2float_rn(t.amount)));
    put(27, t.tx_type == IGS_TX_DEPOSIT ? 1.f : 0.f);
    put(28, t.tx_type == IGS_TX_WITHDRAW ? 1.f : 0.f);
    put(29, t.tx_type == IGS_TX_BET ? 1.f : 0.f);
    int k = 30;                                    // north-star extension block, idx 30..41
#pragma unroll
    for (int ty = 0; ty < 2; ++ty)
#pragma unroll
        for (int w = 0; w < 3; ++w) {
            put(k++, __uint2float_rn(v.win[ty][w].count));
            put(k++, __ll2float_rn(v.win[ty][w].sum));
        }
    if (!kCompact) {
#pragma unroll
        for (int i = 42; i < IGS_NUM_FEATURES; ++i) put(i, 0.f);
    }
    r->amount = t.amount;
    r->total_deposits = pl.r.total_deposits;
    r->total_withdrawals = pl.r.total_withdrawals;
    r->tx_count_1m = v.tx_count_1m;
    r->tslt = v.time_since_last_tx;
    r->deposit_count = (unsigned)pl.r.deposit_count;
    r->uniq_dev = uniq_dev;
    r->uniq_ip = uniq_ip;
    r->bits = (tbits & (IGS_BIT_VPN | IGS_BIT_PROXY | IGS_BIT_TOR | IGS_BIT_BLACKLISTED | IGS_BIT_VALID)) |
              (bonus_only ? IGS_BIT_BONUS_ONLY : 0u);
    r->account_age = account_age;
    r->tx_type = t.tx_type;
}

template <int kSrc, bool kLeaves, typename S>
__global__ void __launch_bounds__(S::kThreads, S::kCtasPerSm)
k_score(const __grid_constant__ ScoreParams p) {
    constexpr int kWarps = S::kWarps, kTileTx = S::kTileTx, kChunkTrees = S::kChunkTrees,
                  kChunkBytes = S::kChunkBytes;
    extern __shared__ __align__(128) uint8_t smem[];
    float* feat_tiles = reinterpret_cast<float*>(smem);
    uint8_t* chunk_buf = smem + kWarps * kFeatTileBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(chunk_buf + 2 * kChunkBytes);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar0 = smem_u32(bars);
    if (threadIdx.x == 0) {
        mbar_init(bar0, 1);
        mbar_init(bar0 + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int model_kind = p.cfg.model_kind;
    const unsigned n_chunks = (p.n_trees + kChunkTrees - 1) / kChunkTrees;
    const bool use_gbdt = model_kind == IGS_MODEL_GBDT && n_chunks > 0;
    const unsigned long long n_tiles = (p.n + kTileTx - 1) / kTileTx;
    const unsigned long long my_tiles =
        n_tiles > blockIdx.x ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const unsigned long long total_chunks = use_gbdt ? my_tiles * n_chunks : 0;

    auto issue = [&](unsigned long long g) {   // g-th chunk of this CTA's sequence; thread 0 only
        const unsigned c = (unsigned)(g % n_chunks);
        const unsigned b = (unsigned)(g & 1);
        const unsigned trees = min((unsigned)kChunkTrees, p.n_trees - c * kChunkTrees);
        const uint32_t bytes = trees * IGS_DEV_TREE_BYTES;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(bar0 + 8 * b, bytes);
        bulk_g2s(smem_u32(chunk_buf + b * kChunkBytes), p.model + (size_t)c * kChunkBytes, bytes,
                 bar0 + 8 * b);
    };
    if (threadIdx.x == 0) {
        if (total_chunks > 0) issue(0);
        if (total_chunks > 1) issue(1);
    }

    float* my_tile = feat_tiles + warp * kFeatTileFloats;
    const uint32_t tile_lane_addr = smem_u32(my_tile + lane);
    const uint32_t chunk_addr = smem_u32(chunk_buf);
    unsigned long long g = 0;

    for (unsigned long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const unsigned long long slot = tile * kTileTx + threadIdx.x;
        const bool in_range = slot < p.n;
        // path B may visit the batch in account order: neighbouring lanes then gather neighbouring (often the
        // same) index records.  Every read and write of the row goes through the permutation.
        const unsigned long long row = (kSrc == kSrcTx && p.perm != nullptr && in_range) ? (uint32_t)__ldg(p.perm + slot) : slot;

        RuleIn ri;
        if (kSrc == kSrcFeatures) {
            // ---- stage this warp's 32 feature rows, transposed to [feature][lane] ------------
            if (model_kind != IGS_MODEL_NONE) {
                const float4* src = reinterpret_cast<const float4*>(p.features + row * IGS_NUM_FEATURES);
#pragma unroll
                for (int q = 0; q < IGS_NUM_FEATURES / 4; ++q) {
                    float4 v = in_range ? __ldg(src + q) : make_float4(0.f, 0.f, 0.f, 0.f);
                    my_tile[(4 * q + 0) * 32 + lane] = v.x;
                    my_tile[(4 * q + 1) * 32 + lane] = v.y;
                    my_tile[(4 * q + 2) * 32 + lane] = v.z;
                    my_tile[(4 * q + 3) * 32 + lane] = v.w;
                }
            }
            // ---- rule sideband ---------------------------------------------------------------
            igs_rule_row rr;
            uint4* dst = reinterpret_cast<uint4*>(&rr);
            const uint4* src = reinterpret_cast<const uint4*>(p.rows + row);
            const uint4 z = make_uint4(0, 0, 0, 0);
            dst[0] = in_range ? __ldg(src + 0) : z;
            dst[1] = in_range ? __ldg(src + 1) : z;
            dst[2] = in_range ? __ldg(src + 2) : z;
            ri.amount = rr.amount; ri.total_deposits = rr.total_deposits;
            ri.total_withdrawals = rr.total_withdrawals; ri.tx_count_1m = rr.tx_count_1m;
            ri.tslt = rr.time_since_last_tx; ri.deposit_count = rr.deposit_count;
            ri.uniq_dev = rr.unique_devices_24h; ri.uniq_ip = rr.unique_ips_24h; ri.bits = rr.bits;
            ri.account_age = rr.account_age_days; ri.tx_type = rr.tx_type;
        } else {
            assemble_tx(p, row, in_range, my_tile + lane, &ri);
            if (p.features_out && in_range) {
                float4* dst = reinterpret_cast<float4*>(p.features_out + row * IGS_NUM_FEATURES);
#pragma unroll
                for (int q = 0; q < IGS_NUM_FEATURES / 4; ++q)
                    dst[q] = make_float4(my_tile[(4 * q + 0) * 32 + lane], my_tile[(4 * q + 1) * 32 + lane],
                                         my_tile[(4 * q + 2) * 32 + lane], my_tile[(4 * q + 3) * 32 + lane]);
            }
        }
        __syncwarp();

        // ---- flags + rule score (applyRules) --------------------------------------------------
        unsigned flags;
        const int rule = igs_apply_rules(p.cfg, ri.amount, ri.tx_type, ri.tx_count_1m, ri.account_age,
                                         ri.uniq_dev, ri.uniq_ip, ri.bits, ri.tslt, ri.deposit_count,
                                         ri.total_deposits, ri.total_withdrawals, &flags);
        const bool valid = (ri.bits & IGS_BIT_VALID) != 0;

        // ---- model ----------------------------------------------------------------------------
        double ml = 0.0;
        if (model_kind == IGS_MODEL_MOCK) {
            ml = mock_predict(my_tile + lane);
        } else if (model_kind == IGS_MODEL_GBDT) {
            float margin = p.base_score;
            uint8_t* leaf_row = kLeaves && in_range ? p.leaf_idx + row * p.n_trees : nullptr;
            for (unsigned c = 0; c < n_chunks; ++c, ++g) {
                const unsigned b = (unsigned)(g & 1);
                const int trees = (int)min((unsigned)kChunkTrees, p.n_trees - c * kChunkTrees);
                mbar_wait(bar0 + 8 * b, (uint32_t)((g >> 1) & 1));
                if (kLeaves) {
                    uint8_t tmp[kChunkTrees];
                    walk_chunk<true, S::kIlp>(chunk_addr + b * kChunkBytes, trees, tile_lane_addr, &margin, tmp);
                    if (leaf_row)
                        for (int t = 0; t < trees; ++t) leaf_row[c * kChunkTrees + t] = valid ? tmp[t] : 0;
                } else {
                    walk_chunk<false, S::kIlp>(chunk_addr + b * kChunkBytes, trees, tile_lane_addr, &margin, nullptr);
                }
                __syncthreads();   // every lane is done with buffer b
                if (threadIdx.x == 0 && g + 2 < total_chunks) issue(g + 2);
            }
            float y = p.post == IGS_POST_LOGISTIC ? igs_sigmoid(margin) : margin;
            if (y < 0.f) y = 0.f;                                   // ml/onnx_model.go:247-252
            if (y > 1.f) y = 1.f;
            ml = (double)y;                                         // float64(result) :254
        }
        if (model_kind != IGS_MODEL_NONE && !valid) ml = 0.5;      // scoring/engine.go:279-282

        // ---- blend + action -------------------------------------------------------------------
        if (in_range) {
            igs_result r;
            igs_blend(p.cfg, rule, flags, ml, model_kind != IGS_MODEL_NONE, &r);
            p.out[row] = r;
        }
        __syncwarp();   // tile is rewritten by the next iteration
    }
}

// ------------------------------------------------------------------------------------------------------------
// Warp-specialised fused kernel (path B, large batches).
// In k_score every warp alternates between two phases with opposite needs: the lookup + assembly of its 32
// transactions (chains of dependent global loads, nothing for the shared-memory pipe to do) and the walk of 500
// trees (shared-memory pipe, no global traffic).  Here the phases run on different warps of ONE persistent CTA per
// SM: kWalkers warps only walk, kProducers warps only look up and assemble — into the OTHER half of each walker's
// double-buffered feature tile — so that the walkers never wait for DRAM and the lookups need no visit-order sort
// to be cheap.  Hand-over per (walker, buffer) through two mbarriers (full / empty); the tree chunks stream through
// shared memory exactly as in k_score, the chunk hand-over being a named barrier over the walker warps only.
// Shared memory: 16 walkers x 2 x 40 rows x 128 B of feature tiles (the 25 features that are constant zero on this
// path share one row: igs_ws_row; the engine keeps a second model image with the feature indices replaced by row
// numbers) + two 10-tree chunk buffers + the rule sideband = 215 KB.
// Registers: 640 threads leave ~100 registers per thread, which the five independent tree walks of a walker need;
// more producer warps would feed the walkers better but push the walk into spills (see launch_fused).
// ------------------------------------------------------------------------------------------------------------
constexpr int kWsRows = IGS_WS_ROWS;
template <int W, int P, int C, int I, int RW = 0, int RP = 0>
struct WsShape {
    static constexpr int kWalkers = W, kProducers = P, kIlp = I, kChunkTrees = C;
    // setmaxnreg targets (0 = leave as compiled).  A CTA never holds more registers than it was launched with: the
    // walkers can only take what the producers have given back (an inc beyond that waits forever).
    static constexpr int kRegsWalker = RW, kRegsProducer = RP;
    static constexpr int kRegsLaunch = (65536 / ((W + P) * 32)) / 8 * 8;
    static_assert(RW == 0 || (W % 4 == 0 && P % 4 == 0), "setmaxnreg works on warpgroups of four warps");
    static_assert(RW == 0 || W * (RW - kRegsLaunch) <= P * (kRegsLaunch - RP), "the walkers' increase must fit the producers' decrease");
    static constexpr int kThreads = (W + P) * 32;
    static constexpr int kTileTx = W * 32;
    static constexpr int kTileBytes = kWsRows * 32 * 4;                 // 5 120
    static constexpr int kChunkBytes = C * IGS_DEV_TREE_BYTES;
    static constexpr int kSideBytes = W * 2 * 32 * 8;                   // {rule score, flags | valid << 31} per lane
    static constexpr int kBarBytes = 16 + W * 2 * 2 * 8;                // 2 chunk barriers + full / empty per (walker, buffer)
    static constexpr int kSmemBytes = W * 2 * kTileBytes + 2 * kChunkBytes + kSideBytes + kBarBytes + 64;
    static_assert(C % I == 0, "chunk must be a whole number of ILP groups");
    static_assert(kSmemBytes <= 227 * 1024, "smem budget");
};

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// bounded wait: a hand-over that never comes is a bug, and a trap beats a hung GPU
__device__ __forceinline__ void mbar_wait_bounded(uint32_t bar, uint32_t parity) {
    uint32_t ok, spins = 0;
    do {
        asm volatile(
            "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            " selp.u32 %0, 1, 0, p;\n}\n"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
        if (!ok && ++spins > (1u << 26)) __trap();
    } while (!ok);
}

// The batch is being copied in while this kernel runs (igs_tx_arrival): wait until the chunk that holds the warp's 32
// records is complete.  The copy stream bumps *tx_ready after each chunk; the records were never read before by this
// SM, so the plain loads that follow see what the copy engine wrote.  Bounded: a copy that never comes traps.
__device__ __forceinline__ void tx_wait_arrival(const ScoreParams& p, unsigned long long first_row, int lane) {
    const unsigned long long last = min(first_row + 31ull, p.n - 1ull);
    const uint32_t need = (uint32_t)(last / p.tx_chunk_rows) + 1u;
    if (lane == 0) {
        uint32_t have, spins = 0;
        for (;;) {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(have) : "l"(p.tx_ready) : "memory");
            if (have >= need) break;
            __nanosleep(256);
            if (++spins > (1u << 25)) __trap();      // ~10 s
        }
    }
    __syncwarp();
}

template <typename S>
__global__ void __launch_bounds__(S::kThreads, 1)
k_score_ws(const __grid_constant__ ScoreParams p) {
    constexpr int W = S::kWalkers, P = S::kProducers, kTileTx = S::kTileTx, kChunkTrees = S::kChunkTrees,
                  kChunkBytes = S::kChunkBytes, kTileBytes = S::kTileBytes;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* tiles = smem;                                              // [W][2][kWsRows][32] f32
    uint8_t* chunk_buf = tiles + W * 2 * kTileBytes;
    uint2* side = reinterpret_cast<uint2*>(chunk_buf + 2 * kChunkBytes);    // [W][2][32]
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(side) + S::kSideBytes);
    const uint32_t bar_chunk = smem_u32(bars);                          // [2]
    const uint32_t bar_full = bar_chunk + 16;                           // [W][2]
    const uint32_t bar_empty = bar_full + W * 2 * 8;                    // [W][2]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(bar_chunk, 1);
        mbar_init(bar_chunk + 8, 1);
        for (int i = 0; i < W * 2; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int model_kind = p.cfg.model_kind;
    const unsigned n_chunks = (p.n_trees + kChunkTrees - 1) / kChunkTrees;
    const bool use_gbdt = model_kind == IGS_MODEL_GBDT && n_chunks > 0;
    const unsigned long long n_tiles = (p.n + kTileTx - 1) / kTileTx;
    const unsigned long long my_tiles = n_tiles > blockIdx.x ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

    if (warp >= W) {
        // ---------------- producers: lookup + assembly + rules for the walkers w = pw, pw + P, ... ----------------
        if (S::kRegsProducer) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(S::kRegsProducer ? S::kRegsProducer : 24));
        const int pw = warp - W;
        for (unsigned long long k = 0; k < my_tiles; ++k) {
            const unsigned long long tile = blockIdx.x + k * gridDim.x;
            const unsigned b = (unsigned)(k & 1);
            for (int w = pw; w < W; w += P) {
                const unsigned long long slot = tile * kTileTx + (unsigned)w * 32 + lane;
                const bool in_range = slot < p.n;
                const unsigned long long row = (p.perm != nullptr && in_range) ? (uint32_t)__ldg(p.perm + slot) : slot;
                mbar_wait_bounded(bar_empty + 8 * (w * 2 + b), (uint32_t)(((k >> 1) & 1) ^ 1));   // the walker is done with it
                if (p.tx_ready != nullptr) tx_wait_arrival(p, tile * kTileTx + (unsigned)w * 32, lane);
                float* tile_lane = reinterpret_cast<float*>(tiles + (size_t)(w * 2 + b) * kTileBytes) + lane;
                RuleIn ri;
                assemble_tx<true>(p, row, in_range, tile_lane, &ri);
                unsigned flags;
                const int rule = igs_apply_rules(p.cfg, ri.amount, ri.tx_type, ri.tx_count_1m, ri.account_age, ri.uniq_dev,
                                                 ri.uniq_ip, ri.bits, ri.tslt, ri.deposit_count, ri.total_deposits,
                                                 ri.total_withdrawals, &flags);
                side[(w * 2 + b) * 32 + lane] = make_uint2((uint32_t)rule, flags | ((ri.bits & IGS_BIT_VALID) ? 0x80000000u : 0u));
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_full + 8 * (w * 2 + b));
            }
        }
        return;
    }

    // ---------------- walkers -------------------------------------------------------------------------------------
    if (S::kRegsWalker) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(S::kRegsWalker ? S::kRegsWalker : 24));
    const unsigned long long total_chunks = use_gbdt ? my_tiles * n_chunks : 0;
    auto issue = [&](unsigned long long g) {   // g-th chunk of this CTA's sequence; thread 0 only
        const unsigned c = (unsigned)(g % n_chunks);
        const unsigned b = (unsigned)(g & 1);
        const unsigned trees = min((unsigned)kChunkTrees, p.n_trees - c * kChunkTrees);
        const uint32_t bytes = trees * IGS_DEV_TREE_BYTES;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(bar_chunk + 8 * b, bytes);
        bulk_g2s(smem_u32(chunk_buf + b * kChunkBytes), p.model + (size_t)c * kChunkBytes, bytes, bar_chunk + 8 * b);
    };
    if (threadIdx.x == 0) {
        if (total_chunks > 0) issue(0);
        if (total_chunks > 1) issue(1);
    }
    const uint32_t chunk_addr = smem_u32(chunk_buf);
    unsigned long long g = 0;
    for (unsigned long long k = 0; k < my_tiles; ++k) {
        const unsigned long long tile = blockIdx.x + k * gridDim.x;
        const unsigned b = (unsigned)(k & 1);
        const unsigned long long slot = tile * kTileTx + (unsigned)warp * 32 + lane;
        const bool in_range = slot < p.n;
        const unsigned long long row = (p.perm != nullptr && in_range) ? (uint32_t)__ldg(p.perm + slot) : slot;
        mbar_wait_bounded(bar_full + 8 * (warp * 2 + b), (uint32_t)((k >> 1) & 1));       // the producer has filled it
        const float* my_tile = reinterpret_cast<const float*>(tiles + (size_t)(warp * 2 + b) * kTileBytes);
        const uint32_t tile_lane_addr = smem_u32(my_tile + lane);
        const uint2 sd = side[(warp * 2 + b) * 32 + lane];
        const int rule = (int)sd.x;
        const unsigned flags = sd.y & 0x7FFFFFFFu;
        const bool valid = (sd.y >> 31) != 0;
        double ml = 0.0;
        if (model_kind == IGS_MODEL_MOCK) {
            ml = mock_predict<true>(my_tile + lane);
        } else if (model_kind == IGS_MODEL_GBDT) {
            float margin = p.base_score;
            for (unsigned c = 0; c < n_chunks; ++c, ++g) {
                const unsigned cb = (unsigned)(g & 1);
                const int trees = (int)min((unsigned)kChunkTrees, p.n_trees - c * kChunkTrees);
                mbar_wait(bar_chunk + 8 * cb, (uint32_t)((g >> 1) & 1));
                walk_chunk<false, S::kIlp>(chunk_addr + cb * kChunkBytes, trees, tile_lane_addr, &margin, nullptr);
                asm volatile("bar.sync 1, %0;" ::"n"(W * 32) : "memory");   // every walker is done with buffer cb
                if (threadIdx.x == 0 && g + 2 < total_chunks) issue(g + 2);
            }
            float y = p.post == IGS_POST_LOGISTIC ? igs_sigmoid(margin) : margin;
            if (y < 0.f) y = 0.f;                                   // ml/onnx_model.go:247-252
            if (y > 1.f) y = 1.f;
            ml = (double)y;                                         // float64(result) :254
        }
        if (model_kind != IGS_MODEL_NONE && !valid) ml = 0.5;      // scoring/engine.go:279-282
        if (in_range) {
            igs_result r;
            igs_blend(p.cfg, rule, flags, ml, model_kind != IGS_MODEL_NONE, &r);
            p.out[row] = r;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_empty + 8 * (warp * 2 + b));     // the producer may refill this half
    }
}

}  // namespace

template <typename S>
static int launch_shape(igs_engine* e, int src, const ScoreParams& p, bool leaves, cudaStream_t stream) {
    void (*kern)(const ScoreParams) = nullptr;
    if (src == kSrcFeatures) kern = leaves ? k_score<kSrcFeatures, true, S> : k_score<kSrcFeatures, false, S>;
    else kern = leaves ? k_score<kSrcTx, true, S> : k_score<kSrcTx, false, S>;
    // per-device attribute; a few hundred ns, so simply set on every launch
    IGS_CUDA(e, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kSmemBytes));
    const unsigned long long n_tiles = (p.n + S::kTileTx - 1) / S::kTileTx;
    const unsigned long long max_ctas = (unsigned long long)e->sm_count * S::kCtasPerSm;
    const unsigned grid = (unsigned)(n_tiles < max_ctas ? n_tiles : max_ctas);
    kern<<<grid, S::kThreads, S::kSmemBytes, stream>>>(p);
    IGS_LAUNCHED(e);
    return IGS_OK;
}

static int launch(igs_engine* e, int src, const ScoreParams& p, bool leaves, cudaStream_t stream) {
    static const int variant = getenv("IGS_SCORE_SHAPE") ? atoi(getenv("IGS_SCORE_SHAPE")) : 0;   // tuning knob
    switch (variant) {          // alternatives kept for tuning runs (profiles/README.md lists what they measured)
    case 1: return launch_shape<Shape<12, 20>>(e, src, p, leaves, stream);
    case 2: return launch_shape<Shape<16, 10>>(e, src, p, leaves, stream);
    case 3: return launch_shape<Shape<8, 25>>(e, src, p, leaves, stream);
    case 4: return launch_shape<Shape<16, 14, 7>>(e, src, p, leaves, stream);
    case 5: return launch_shape<Shape<6, 5, 5, 3>>(e, src, p, leaves, stream);     // three small CTAs per SM
    default:
        // latency-sized batches: if the 12-warp shape would need a second, nearly empty round of tiles while
        // the 16-warp shape covers the batch in one round, take the latter
        if (src == kSrcFeatures && p.n > (unsigned long long)e->sm_count * Shape<12, 20>::kTileTx &&
            p.n <= (unsigned long long)e->sm_count * Shape<16, 14, 7>::kTileTx)
            return launch_shape<Shape<16, 14, 7>>(e, src, p, leaves, stream);
        // fused path: one round of 16-warp tiles when the batch fits in it (lowest latency), else two 8-warp CTAs
        // per SM (highest throughput)
        if (src == kSrcTx)
            return p.n <= (unsigned long long)e->sm_count * Shape<16, 14, 7>::kTileTx
                       ? launch_shape<Shape<16, 14, 7>>(e, src, p, leaves, stream)
                       : launch_shape<Shape<8, 10, 5, 2>>(e, src, p, leaves, stream);
        return launch_shape<Shape<12, 20>>(e, src, p, leaves, stream);
    }
}

static int check_model(igs_engine* e, const igs_config* cfg) {
    IGS_REQUIRE(e, cfg->model_kind >= IGS_MODEL_NONE && cfg->model_kind <= IGS_MODEL_GBDT,
                "cfg.model_kind");
    if (cfg->model_kind == IGS_MODEL_GBDT && !e->has_gbdt) {
        igs_set_error(e, "model_kind = IGS_MODEL_GBDT but no ensemble was loaded (igs_load_gbdt)");
        return IGS_ERR_STATE;
    }
    return IGS_OK;
}

static void fill_common(igs_engine* e, const igs_config* cfg, ScoreParams* p) {
    memset(p, 0, sizeof(*p));
    p->cfg = *cfg;
    p->model = e->d_model;
    p->model_ws = e->d_model_ws;
    p->n_trees = e->has_gbdt ? e->n_trees : 0;
    p->base_score = e->base_score;
    p->post = e->post_transform;
}

int igs_launch_score_features(igs_engine* e, const igs_config* cfg, const float* d_features,
                              const igs_rule_row* d_rows, uint64_t n, igs_result* d_out,
                              uint8_t* d_leaf_idx, cudaStream_t stream) {
    if (n == 0) return IGS_OK;
    int rc = check_model(e, cfg);
    if (rc != IGS_OK) return rc;
    IGS_REQUIRE(e, d_rows && d_out, "rows / out pointer is null");
    IGS_REQUIRE(e, cfg->model_kind == IGS_MODEL_NONE || d_features, "features pointer is null");
    ScoreParams p;
    fill_common(e, cfg, &p);
    p.features = d_features;
    p.rows = d_rows;
    p.out = d_out;
    p.leaf_idx = d_leaf_idx;
    p.n = n;
    const bool leaves = d_leaf_idx != nullptr && cfg->model_kind == IGS_MODEL_GBDT;
    return launch(e, kSrcFeatures, p, leaves, stream);
}

// ---- account-ordered visiting of large fused batches -----------------------------------------------------
// The index gathers of the fused path are bound by random 32-byte sector reads.  Visiting the batch in
// account order makes neighbouring lanes read neighbouring records (the same ones for a hot account): one or
// three radix passes over (account, row) cost less than they save once the batch is large (1M tx: 1.85 -> 1.76
// ms including the sort; one pass — locality without adjacency — gains nothing).  The scratch is stream-ordered, from
// the engine's own pool (igs_engine::scratch_pool), so concurrent calls on different streams do not share it.
// (Host-buffer calls arrive here in 131 072-row chunks, below this threshold: sorting them was measured and costs
// more than it saves end to end — 2.03 -> 2.17 ms per 1M tx through igs_score_batch.)
constexpr unsigned long long kVisitSortMin = 262144;
// (A counting sort with an atomic cursor per account — the visit order need not be stable — was measured: the hot
// account's ~30 000 same-address atomics per kernel make it slower than the radix passes, 1.67 -> 1.86 ms per 1M tx;
// grouping by buckets of 16 accounts instead of by account: 1.95 ms, worse than no grouping.)
// key = [account (top bits) | row]: sorted on the account bits by the one-sweep radix sort (onesweep.cuh); the row
// rides in the low word and is what the score kernel reads back.  The key pass also fills the digit histograms.
__global__ void __launch_bounds__(kOsThreads)
k_visit_keys(const igs_tx* __restrict__ tx, uint32_t n, OsPlan plan, OsScratch sc, uint64_t* __restrict__ keys) {
    extern __shared__ uint32_t vk_sh_hist[];
    __shared__ OsPlan sh_plan;
    if (threadIdx.x == 0) sh_plan = plan;
    os_hist_clear(vk_sh_hist, plan.passes);
    __syncthreads();
    for (uint32_t i = blockIdx.x * kOsThreads + threadIdx.x; i < n; i += gridDim.x * kOsThreads) {
        const uint64_t k = ((uint64_t)__ldg(&tx[i].account) << 32) | i;
        keys[i] = k;
        os_hist_accumulate(vk_sh_hist, sh_plan, k);
    }
    __syncthreads();
    os_hist_flush(vk_sh_hist, sh_plan, sc);
}
static int visit_order(igs_engine* e, const igs_tx* d_tx, uint32_t n, uint32_t n_accounts, int passes, cudaStream_t stream,
                       void** scratch, const uint64_t** perm) {
    int acct_bits = 1;
    while (acct_bits < 32 && (n_accounts >> acct_bits)) ++acct_bits;
    // the top `passes` digits of the account: locality needs the accounts grouped, not fully ordered
    const int begin = 32 + std::max(0, acct_bits - passes * 8), end = 32 + acct_bits;
    const OsPlan plan = os_plan(begin, end);
    const size_t b_keys = ((size_t)n * 8 + 255) & ~(size_t)255;
    const size_t b_os = (OsScratch::words(n, plan.passes) * 4 + 255) & ~(size_t)255;
    char* base = nullptr;
    IGS_CUDA(e, cudaMallocFromPoolAsync((void**)&base, 2 * b_keys + b_os, e->scratch_pool, stream));
    *scratch = base;
    uint64_t* keys[2] = {(uint64_t*)base, (uint64_t*)(base + b_keys)};
    OsScratch sc;
    int rc = os_prepare(e, &sc, (uint32_t*)(base + 2 * b_keys), n, plan, stream);
    if (rc != IGS_OK) return rc;
    IGS_CUDA(e, cudaFuncSetAttribute(k_visit_keys, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)os_hist_smem_bytes(plan.passes)));
    k_visit_keys<<<std::min<uint32_t>((n + kOsThreads - 1) / kOsThreads, (uint32_t)e->sm_count * 8), kOsThreads,
                   os_hist_smem_bytes(plan.passes), stream>>>(d_tx, n, plan, sc, keys[0]);
    IGS_LAUNCHED(e);
    int cur = 0;
    rc = os_sort_passes(e, keys, n, plan, sc, stream, &cur);
    if (rc != IGS_OK) return rc;
    *perm = keys[cur];
    return IGS_OK;
}

// Large fused batches that want nothing but the results go through the warp-specialised kernel; the rest (latency-
// sized batches, calls that also return the assembled features or the velocity blocks) through k_score.
template <typename S>
static int launch_ws(igs_engine* e, const ScoreParams& p, cudaStream_t stream) {
    auto kern = k_score_ws<S>;
    IGS_CUDA(e, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kSmemBytes));
    const unsigned long long n_tiles = (p.n + S::kTileTx - 1) / S::kTileTx;
    const unsigned grid = (unsigned)(n_tiles < (unsigned long long)e->sm_count ? n_tiles : (unsigned long long)e->sm_count);
    kern<<<grid, S::kThreads, S::kSmemBytes, stream>>>(p);
    IGS_LAUNCHED(e);
    return IGS_OK;
}
static int launch_fused(igs_engine* e, ScoreParams& p, cudaStream_t stream) {
    static const int ws = getenv("IGS_SCORE_WS") ? atoi(getenv("IGS_SCORE_WS")) : 1;      // tuning knob: 0 = never
    // at least four tiles per SM: with fewer the producers' head start on the first tile and the drain of the last
    // are not amortised (the 131 072-row chunks of a host-buffer call were measured slower through this kernel)
    const bool eligible = ws != 0 && p.features_out == nullptr && p.velocity_out == nullptr &&
                          p.n >= 4ull * (unsigned long long)e->sm_count * Shape<16, 14, 7>::kTileTx &&
                          (p.cfg.model_kind != IGS_MODEL_GBDT || e->d_model_ws != nullptr);
    if (!eligible) {
        IGS_REQUIRE(e, p.tx_ready == nullptr, "a batch that is still arriving needs the warp-specialised kernel");
        return launch(e, kSrcTx, p, false, stream);
    }
    p.model = p.model_ws ? p.model_ws : p.model;
    // 16 walkers + 4 producers (640 threads, ~100 registers each).  Measured alternatives, 1M tx against 500k players
    // and a 5M-event index, ms per batch with / without the visit-order sort (k_score: 1.757 / 1.959):
    //   16 + 4: 1.671 / 1.749     16 + 6: 1.725 / 1.797     16 + 8: 1.774 / 1.908 (80 registers: the walk spills)
    //   12 + 8: 1.695 / 1.722
    //   16 + 8 re-balanced with setmaxnreg (walkers 96 / producers 48 registers): 1.731 / 1.874 — the producers spill
    return launch_ws<WsShape<16, 4, 10, 5>>(e, p, stream);
}

// batches the warp-specialised kernel takes (launch_fused) — the host-buffer call reads such batches straight from
// pinned host memory instead of staging them
bool igs_fused_direct_eligible(const igs_engine* e, uint64_t n) {
    static const int ws = getenv("IGS_SCORE_WS") ? atoi(getenv("IGS_SCORE_WS")) : 1;
    static const int direct = getenv("IGS_SCORE_DIRECT") ? atoi(getenv("IGS_SCORE_DIRECT")) : 1;   // tuning knob: 0 = always stage
    return ws != 0 && direct != 0 && n >= 4ull * (unsigned long long)e->sm_count * Shape<16, 14, 7>::kTileTx &&
           (!e->has_gbdt || e->d_model_ws != nullptr);
}

int igs_launch_score_batch(igs_engine* e, const igs_config* cfg, const igs_tx* d_tx, const igs_tx_ids* d_ids,
                           uint64_t n, igs_result* d_out, float* d_features_out, igs_velocity* d_velocity_out,
                           cudaStream_t stream, const igs_tx_arrival* arrival, bool tx_over_pcie) {
    if (n == 0) return IGS_OK;
    int rc = check_model(e, cfg);
    if (rc != IGS_OK) return rc;
    ScoreParams p;
    fill_common(e, cfg, &p);
    p.tx = d_tx;
    p.players = e->d_players;
    p.n_players = e->d_players ? e->n_players : 0;
    igs_history_reader rd(e);          // the published base + delta; n_accounts = 0 when no history is loaded
    p.ix = rd.view;
    p.features_out = d_features_out;
    p.velocity_out = d_velocity_out;
    p.tx_ids = d_ids;
    igs_blacklist_views(e, p.bl);
    p.uniques_from_index = e->has_events && e->has_event_ids;
    p.out = d_out;
    p.n = n;
    static const int visit_passes = getenv("IGS_VISIT_SORT_PASSES") ? atoi(getenv("IGS_VISIT_SORT_PASSES")) : 3;   // tuning knob (0 = off)
    void* scratch = nullptr;
    if (arrival) {
        IGS_REQUIRE(e, igs_fused_direct_eligible(e, n) && !d_features_out && !d_velocity_out && arrival->chunk_rows > 0,
                    "a batch that is still arriving needs the warp-specialised kernel");
        p.tx_ready = arrival->ready;
        p.tx_chunk_rows = arrival->chunk_rows;
    }
    if (visit_passes > 0 && !tx_over_pcie && !arrival && e->has_events && n >= kVisitSortMin && n < 0xFFFFFFFFull) {
        rc = visit_order(e, d_tx, (uint32_t)n, rd.view.base.n_accounts, std::min(visit_passes, 4), stream, &scratch, &p.perm);
        if (rc != IGS_OK) { if (scratch) cudaFreeAsync(scratch, stream); return rc; }
    }
    rc = launch_fused(e, p, stream);
    if (rc == IGS_OK) rd.done(stream);
    if (scratch) cudaFreeAsync(scratch, stream);
    return rc;
}
