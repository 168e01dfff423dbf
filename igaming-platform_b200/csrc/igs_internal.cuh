// igs_internal.cuh — engine object and helpers shared by the translation units of libigs.so.
// Not part of the ABI (that is include/igs.h).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <mutex>
#include <vector>

#include "igs.h"
#include "velocity.cuh"

// ---------------------------------------------------------------------------------------------
// device-resident model layout
// ---------------------------------------------------------------------------------------------
// One tree = 2384 B, contiguous and 16B-aligned so that a chunk of trees is one cp.async.bulk.
// Levels are packed in pairs so that a walk needs 5 shared-memory wavefronts per two levels instead
// of 6: an even-level node (8 bytes) carries its threshold, its own feature and the features of
// both children; an odd-level node is only a 4-byte threshold (its feature came with the parent).
//   level 0  1 x 8 B  @    0      level 1    2 x 4 B @    8
//   level 2  4 x 8 B  @   16      level 3    8 x 4 B @   48
//   level 4 16 x 8 B  @   80      level 5   32 x 4 B @  208
//   level 6 64 x 8 B  @  336      level 7  128 x 4 B @  848
//   leaves 256 x 4 B  @ 1360
// A level of <= 128 contiguous bytes covers every bank at most once, so levels 0..5 are free of
// bank conflicts whatever nodes the 32 lanes visit.
struct __align__(8) igs_dev_node {
    uint32_t feats;   // own feature * 128 (bits 7..12) | left child's feature << 13 | right child's << 19
    float    thr;
};
__host__ __device__ constexpr int igs_dev_level_off(int d) {   // 0, 8, 16, 48, 80, 208, 336, 848
    int off = 0;
    for (int l = 0; l < d; ++l) off += (1 << l) * ((l & 1) ? 4 : 8);
    return off;
}
static_assert(igs_dev_level_off(7) == 848 && igs_dev_level_off(8) == 1360, "tree image layout");
constexpr int IGS_DEV_LEAF_OFF = 1360;
// On the fused path (assembly on the device) 39 of the 64 model inputs are live: 7 (IPCountryChanges), 8 (DeviceAge)
// and 22 (DisposableEmail) are never set by the reference (scoring/engine.go:326-417) and 42..63 are spare.  The
// warp-specialised kernel keeps one shared zero row for all of them: feature index -> tile row.
constexpr int IGS_WS_ROWS = 40;
__host__ __device__ constexpr int igs_ws_row(int f) {
    return f < 7 ? f : (f >= 9 && f <= 21) ? f - 2 : (f >= 23 && f <= 41) ? f - 3 : IGS_WS_ROWS - 1;
}
static_assert(igs_ws_row(6) == 6 && igs_ws_row(9) == 7 && igs_ws_row(21) == 19 && igs_ws_row(23) == 20 &&
              igs_ws_row(41) == 38 && igs_ws_row(7) == 39 && igs_ws_row(63) == 39, "row map");
constexpr int IGS_DEV_TREE_BYTES = 1360 + 256 * 4;  // 2384

// ---------------------------------------------------------------------------------------------
// per-slot staging (device buffers that mirror the caller's host buffers)
// ---------------------------------------------------------------------------------------------
struct igs_buf {
    void*  p = nullptr;
    size_t cap = 0;
};

constexpr int IGS_PIPE_STREAMS = 3;   // chunk pipeline depth of a host-buffer call

struct igs_slot {
    cudaStream_t stream = nullptr;                    // the slot's stream: igs_wait() syncs this one
    cudaStream_t pipe[IGS_PIPE_STREAMS] = {nullptr};  // workers of the chunk pipeline
    cudaEvent_t  ev_fork = nullptr;                   // slot stream -> workers
    cudaEvent_t  ev_join[IGS_PIPE_STREAMS] = {nullptr};  // workers -> slot stream
    igs_buf      in0, in1, out0, aux0, aux1;
    uint32_t*    d_ready = nullptr;                   // arrival counter of a batch that is copied in while its kernel runs
};

// A transaction batch that is still arriving in device memory while the kernel that scores it runs: rows
// [c * chunk_rows, (c + 1) * chunk_rows) are complete once *ready > c (written by the copy stream after each chunk).
struct igs_tx_arrival {
    const uint32_t* ready;
    uint32_t chunk_rows;
};
constexpr int IGS_ARRIVAL_CHUNKS = 32;

// one blacklist set: open-addressing table of 64-bit ids, at most half full
struct igs_idset {
    uint64_t* slots = nullptr;
    uint64_t  cap = 0;        // power of two, 0 = no table
    uint64_t  count = 0;      // distinct ids stored
};

// event history index (see velocity.cuh / k_velocity.cu)
struct igs_prefix_rec;
struct igs_tail_rec;
struct igs_event_index {
    uint64_t n = 0;
    uint32_t n_accounts = 0;
    uint64_t cap_events = 0;          // capacity of ts / seq / pre / tail (grow-only, 1.5x)
    uint64_t cap_accounts = 0;        // capacity of seg
    uint32_t* seg = nullptr;          // [n_accounts + 1] first sorted position of each account
    uint32_t* ts = nullptr;           // [n] sorted by (account, ts, arrival)
    uint32_t* seq = nullptr;          // [n] arrival index of the sorted event
    igs_prefix_rec* pre = nullptr;    // [n + 1] exclusive prefix columns
    igs_tail_rec* tail = nullptr;     // [n] inclusive chain columns
    // kernels launched while this set of arrays was the published one: one event each, recorded on the
    // launching stream right after the launch.  The ingest stream waits for them before it overwrites the
    // arrays (igs_history_reader in k_velocity.cu).
    std::vector<cudaEvent_t> readers;
};

struct igs_engine {
    int  device = 0;
    int  sm_count = 0;
    char err[512] = {0};
    std::mutex mu;                // guards loads vs. scoring calls
    igs_slot slot[IGS_NUM_SLOTS];
    igs_buf  part_scratch;        // shard-partition counters (k_partition.cu)
    // Stream-ordered scratch of the scoring calls (the visit-order sort of large fused batches): an engine-private
    // pool that keeps what it has allocated (release threshold = max), so that a call's cudaMallocFromPoolAsync /
    // cudaFreeAsync pair costs microseconds instead of a round trip to the driver, and calls on different streams
    // still get disjoint memory.
    cudaMemPool_t scratch_pool = nullptr;
    uint32_t* h_arrival_vals = nullptr;  // pinned: 1, 2, ... IGS_ARRIVAL_CHUNKS (sources of the arrival-counter copies)
    int mlp_max_pairs[3] = {0, 0, 0};   // co-resident CTA pairs of the three k_mlp_gemm pair variants (0 = not asked yet)

    // tree ensemble
    uint8_t* d_model = nullptr;   // n_trees * IGS_DEV_TREE_BYTES
    uint8_t* d_model_ws = nullptr;  // same, feature indices replaced by igs_ws_row (k_score_ws)
    uint32_t n_trees = 0, n_features = 0, post_transform = 0;
    float    base_score = 0.f;
    bool     has_gbdt = false;

    // players
    igs_player_row* d_players = nullptr;
    uint64_t n_players = 0;

    // events: device-resident raw columns in arrival order (igs_load_events + igs_append_events) and
    // the index built from them
    uint32_t* raw_account = nullptr;
    uint32_t* raw_ts = nullptr;
    int64_t*  raw_amount = nullptr;
    uint8_t*  raw_type = nullptr;
    uint64_t* raw_device = nullptr;   // identifier columns, only when the history carries identifiers
    uint64_t* raw_ip = nullptr;
    bool      has_event_ids = false;
    uint64_t  n_raw = 0, cap_raw = 0;
    // The history as scoring sees it: a base index over raw[0, base_n) and a delta index over
    // raw[base_n, n_raw) (empty after a bulk load / compaction).  Each has a spare set of arrays: a build
    // writes the spare, then the two are swapped under view_mu — kernels already launched keep reading the
    // arrays they were launched with.
    igs_event_index ev, ev_spare;     // base
    igs_event_index dv, dv_spare;     // delta
    uint64_t base_n = 0;
    uint64_t delta_min = 65536;       // igs_set_compaction_threshold
    std::mutex view_mu;               // guards ev / dv (pointers, sizes) and the readers lists
    std::vector<cudaEvent_t> free_events;
    cudaStream_t ingest_stream = nullptr;   // every index build runs here
    bool has_events = false;
    std::vector<igs_buf> build_pool;  // scratch of the index build, kept between igs_append_events calls
    int      late_policy = IGS_LATE_REJECT;   // igs_set_late_policy
    uint64_t late_events = 0;                 // late pairs in the resident history (igs_history_info)

    // blacklist sets (idset.cuh / k_ids.cu), indexed by IGS_BL_*
    igs_idset bl[3];

    // linked accounts (k_links.cu): stage scratch + the per-account table built by igs_links_finish
    std::vector<igs_buf> links_pool;
    igs_links_row* d_links_rows = nullptr;
    uint64_t n_links_rows = 0;
    // peer-memory exchange: this shard's receive window, and every shard's window as seen from here
    void*    links_window = nullptr;          // [cursor block][cap_pairs pairs][cap_links links]
    uint64_t links_cap_pairs = 0, links_cap_links = 0;
    void*    links_peer[256] = {nullptr};     // indexed by shard; own entry = links_window
    bool     links_peer_ipc[256] = {false};   // opened through an IPC handle (to be closed)

    // MLP
    void* d_mlp = nullptr;        // see k_ltv.cu
    bool  has_mlp = false;
};

// ---------------------------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------------------------
void igs_set_error(igs_engine* e, const char* fmt, ...);

#define IGS_CUDA(e, call)                                                                  \
    do {                                                                                   \
        cudaError_t st__ = (call);                                                         \
        if (st__ != cudaSuccess) {                                                         \
            igs_set_error((e), "%s failed: %s (%s:%d)", #call, cudaGetErrorString(st__),   \
                          __FILE__, __LINE__);                                             \
            return IGS_ERR_CUDA;                                                           \
        }                                                                                  \
    } while (0)

// After every kernel launch: counts it (igs_kernel_launches: what bench.py reports as gpu_launches) and
// checks the launch status.
extern unsigned long long g_igs_launches;
#define IGS_LAUNCHED(e)                                                   \
    do {                                                                  \
        __atomic_fetch_add(&g_igs_launches, 1ull, __ATOMIC_RELAXED);      \
        IGS_CUDA((e), cudaGetLastError());                                \
    } while (0)

#define IGS_REQUIRE(e, cond, msg)                                  \
    do {                                                           \
        if (!(cond)) {                                             \
            igs_set_error((e), "invalid argument: %s", (msg));     \
            return IGS_ERR_INVALID;                                \
        }                                                          \
    } while (0)

int igs_reserve(igs_engine* e, igs_buf* b, size_t bytes);

// kernel launchers (defined next to their kernels)
int igs_launch_score_features(igs_engine* e, const igs_config* cfg, const float* d_features,
                              const igs_rule_row* d_rows, uint64_t n, igs_result* d_out,
                              uint8_t* d_leaf_idx, cudaStream_t stream);
// tx_over_pcie: d_tx / d_ids / d_out are mapped host memory — read each record once (no visit-order sort)
int igs_launch_score_batch(igs_engine* e, const igs_config* cfg, const igs_tx* d_tx, const igs_tx_ids* d_ids,
                           uint64_t n, igs_result* d_out, float* d_features_out, igs_velocity* d_velocity_out,
                           cudaStream_t stream, const igs_tx_arrival* arrival = nullptr, bool tx_over_pcie = false);
bool igs_fused_direct_eligible(const igs_engine* e, uint64_t n);
int igs_launch_velocity(igs_engine* e, const igs_query* d_q, uint64_t n, igs_velocity* d_out,
                        cudaStream_t stream);
int igs_launch_uniques(igs_engine* e, const igs_query* d_q, uint64_t n, igs_uniques* d_out, cudaStream_t stream);
// ids == nullptr: no identifier columns (load) / events without identifiers (append)
int igs_build_event_index(igs_engine* e, const igs_event_cols* ev, const igs_event_ids* ids, uint64_t n,
                          uint32_t n_accounts);
int igs_append_event_columns(igs_engine* e, const igs_event_cols* ev, const igs_event_ids* ids,
                             uint64_t n);   // append + rebuild
// linked accounts (k_links.cu)
void igs_links_free(igs_engine* e);
// blacklist sets (k_ids.cu)
int igs_idset_load(igs_engine* e, igs_idset* s, const uint64_t* ids, uint64_t n, bool replace);
void igs_idset_free(igs_idset* s);
int igs_launch_blacklist_check(igs_engine* e, const igs_tx_ids* d_ids, uint64_t n, uint8_t* d_out,
                               cudaStream_t stream);
void igs_free_event_index(igs_engine* e);
void igs_free_history_events(igs_engine* e);
// Snapshot of the published history for the kernel(s) of one call.  Holds the engine's view lock from construction
// until done(): take it, launch what reads `view`, call done(stream) — it records, per index level, an event on the
// launching stream that the next build overwriting those arrays will wait for.
struct igs_history_reader {
    igs_engine* e;
    igs_history_view view;
    explicit igs_history_reader(igs_engine* e);
    void done(cudaStream_t st);
    ~igs_history_reader();
    igs_history_reader(const igs_history_reader&) = delete;
    igs_history_reader& operator=(const igs_history_reader&) = delete;
   private:
    bool locked;
};
int igs_evict_event_history(igs_engine* e, uint32_t now, uint64_t* n_evicted);
int igs_prepare_event_ingest(igs_engine* e, uint64_t max_batch);
int igs_mlp_upload(igs_engine* e, const igs_mlp_desc* d);
void igs_mlp_free(igs_engine* e);
int igs_launch_ltv(igs_engine* e, const uint16_t* d_x, const igs_ltv_row* d_rows, uint64_t n,
                   igs_ltv_result* d_out, float* d_ltv_raw, cudaStream_t stream);
int igs_partition_run(igs_engine* e, const igs_tx* d_tx, uint64_t n, uint32_t n_shards, uint32_t* counts,
                      igs_tx* d_tx_out, uint32_t* d_perm, uint32_t* d_offsets, cudaStream_t stream);
size_t igs_partition_scratch_words(uint64_t n, uint32_t n_shards);
int igs_launch_ltv_hidden(igs_engine* e, const uint16_t* d_x, uint64_t n, int layer, uint16_t* d_hidden,
                          cudaStream_t stream);
