// igs_engine.cu — the C ABI of include/igs.h: engine life cycle, loads, host-buffer staging.
// All compute is in the k_*.cu kernels; there is no CPU fallback in this file or anywhere else
// in the library.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <new>
#include <vector>

#include "igs_internal.cuh"

static thread_local char g_create_error[512] = "";
unsigned long long g_igs_launches = 0;

void igs_set_error(igs_engine* e, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(e ? e->err : g_create_error, 512, fmt, ap);
    va_end(ap);
}

int igs_reserve(igs_engine* e, igs_buf* b, size_t bytes) {
    if (bytes <= b->cap) return IGS_OK;
    if (b->p) {
        IGS_CUDA(e, cudaFree(b->p));
        b->p = nullptr;
        b->cap = 0;
    }
    size_t cap = std::max(bytes, (size_t)1 << 20);
    cap = (cap + 255) & ~(size_t)255;
    cudaError_t st = cudaMalloc(&b->p, cap);
    if (st != cudaSuccess) {
        igs_set_error(e, "cudaMalloc(%zu) failed: %s", cap, cudaGetErrorString(st));
        b->p = nullptr;
        return st == cudaErrorMemoryAllocation ? IGS_ERR_NOMEM : IGS_ERR_CUDA;
    }
    b->cap = cap;
    return IGS_OK;
}

static void free_buf(igs_buf* b) {
    if (b->p) cudaFree(b->p);
    b->p = nullptr;
    b->cap = 0;
}

#define IGS_TRY(call)                      \
    do {                                   \
        int rc__ = (call);                 \
        if (rc__ != IGS_OK) return rc__;   \
    } while (0)

#define IGS_ENGINE(e)                                        \
    do {                                                     \
        if (!(e)) {                                          \
            igs_set_error(nullptr, "engine handle is null"); \
            return IGS_ERR_INVALID;                          \
        }                                                    \
        cudaError_t sd__ = cudaSetDevice((e)->device);       \
        if (sd__ != cudaSuccess) {                           \
            igs_set_error((e), "cudaSetDevice(%d) failed: %s", (e)->device, cudaGetErrorString(sd__)); \
            return IGS_ERR_CUDA;                             \
        }                                                    \
    } while (0)

extern "C" {

// ------------------------------------------------------------------------------------------------
// life cycle
// ------------------------------------------------------------------------------------------------
int igs_engine_create(int device, igs_engine** out) {
    if (!out) {
        igs_set_error(nullptr, "out pointer is null");
        return IGS_ERR_INVALID;
    }
    *out = nullptr;
    int count = 0;
    cudaError_t st = cudaGetDeviceCount(&count);
    if (st != cudaSuccess || count == 0) {
        igs_set_error(nullptr, "no CUDA device available (%s); this library has no CPU fallback",
                      st != cudaSuccess ? cudaGetErrorString(st) : "device count is 0");
        return IGS_ERR_CUDA;
    }
    if (device < 0 || device >= count) {
        igs_set_error(nullptr, "device %d out of range (0..%d)", device, count - 1);
        return IGS_ERR_INVALID;
    }
    st = cudaSetDevice(device);
    if (st != cudaSuccess) {
        igs_set_error(nullptr, "cudaSetDevice(%d) failed: %s", device, cudaGetErrorString(st));
        return IGS_ERR_CUDA;
    }
    cudaDeviceProp prop;
    st = cudaGetDeviceProperties(&prop, device);
    if (st != cudaSuccess) {
        igs_set_error(nullptr, "cudaGetDeviceProperties failed: %s", cudaGetErrorString(st));
        return IGS_ERR_CUDA;
    }
    if (prop.major != 10) {
        igs_set_error(nullptr, "device %d is sm_%d%d; this library is built for sm_100a only", device,
                      prop.major, prop.minor);
        return IGS_ERR_CUDA;
    }
    igs_engine* e = new (std::nothrow) igs_engine();
    if (!e) {
        igs_set_error(nullptr, "out of host memory");
        return IGS_ERR_NOMEM;
    }
    e->device = device;
    e->sm_count = prop.multiProcessorCount;
    {
        cudaMemPoolProps props;
        memset(&props, 0, sizeof(props));
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = device;
        unsigned long long keep = ~0ull;
        if (cudaMemPoolCreate(&e->scratch_pool, &props) != cudaSuccess ||
            cudaMemPoolSetAttribute(e->scratch_pool, cudaMemPoolAttrReleaseThreshold, &keep) != cudaSuccess) {
            igs_set_error(nullptr, "memory pool creation failed: %s", cudaGetErrorString(cudaGetLastError()));
            igs_engine_destroy(e);
            return IGS_ERR_CUDA;
        }
    }
    if (cudaStreamCreateWithFlags(&e->ingest_stream, cudaStreamNonBlocking) != cudaSuccess) {
        igs_set_error(nullptr, "stream creation failed: %s", cudaGetErrorString(cudaGetLastError()));
        igs_engine_destroy(e);
        return IGS_ERR_CUDA;
    }
    for (int s = 0; s < IGS_NUM_SLOTS; ++s) {
        igs_slot& sl = e->slot[s];
        bool ok = cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking) == cudaSuccess &&
                  cudaEventCreateWithFlags(&sl.ev_fork, cudaEventDisableTiming) == cudaSuccess;
        for (int k = 0; ok && k < IGS_PIPE_STREAMS; ++k)
            ok = cudaStreamCreateWithFlags(&sl.pipe[k], cudaStreamNonBlocking) == cudaSuccess &&
                 cudaEventCreateWithFlags(&sl.ev_join[k], cudaEventDisableTiming) == cudaSuccess;
        ok = ok && cudaMalloc(&sl.d_ready, 256) == cudaSuccess;
        if (!ok) {
            igs_set_error(nullptr, "stream/event creation failed: %s",
                          cudaGetErrorString(cudaGetLastError()));
            igs_engine_destroy(e);
            return IGS_ERR_CUDA;
        }
    }
    if (cudaHostAlloc(&e->h_arrival_vals, IGS_ARRIVAL_CHUNKS * sizeof(uint32_t), cudaHostAllocDefault) != cudaSuccess) {
        igs_set_error(nullptr, "pinned allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
        igs_engine_destroy(e);
        return IGS_ERR_CUDA;
    }
    for (int c = 0; c < IGS_ARRIVAL_CHUNKS; ++c) e->h_arrival_vals[c] = (uint32_t)c + 1u;
    *out = e;
    return IGS_OK;
}

void igs_engine_destroy(igs_engine* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    cudaDeviceSynchronize();
    for (int s = 0; s < IGS_NUM_SLOTS; ++s) {
        igs_slot& sl = e->slot[s];
        free_buf(&sl.in0); free_buf(&sl.in1); free_buf(&sl.out0); free_buf(&sl.aux0); free_buf(&sl.aux1);
        if (sl.ev_fork) cudaEventDestroy(sl.ev_fork);
        for (int k = 0; k < IGS_PIPE_STREAMS; ++k) {
            if (sl.ev_join[k]) cudaEventDestroy(sl.ev_join[k]);
            if (sl.pipe[k]) cudaStreamDestroy(sl.pipe[k]);
        }
        if (sl.stream) cudaStreamDestroy(sl.stream);
        if (sl.d_ready) cudaFree(sl.d_ready);
    }
    if (e->h_arrival_vals) cudaFreeHost(e->h_arrival_vals);
    free_buf(&e->part_scratch);
    if (e->ingest_stream) cudaStreamDestroy(e->ingest_stream);
    if (e->scratch_pool) cudaMemPoolDestroy(e->scratch_pool);
    if (e->d_model) cudaFree(e->d_model);
    if (e->d_model_ws) cudaFree(e->d_model_ws);
    if (e->d_players) cudaFree(e->d_players);
    igs_free_event_index(e);
    igs_free_history_events(e);
    for (int k = 0; k < 3; ++k) igs_idset_free(&e->bl[k]);
    igs_links_free(e);
    igs_mlp_free(e);
    delete e;
}

const char* igs_last_error(const igs_engine* e) { return e ? e->err : g_create_error; }

void* igs_alloc_pinned(size_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}
void igs_free_pinned(void* p) {
    if (p) cudaFreeHost(p);
}

// ------------------------------------------------------------------------------------------------
// loads
// ------------------------------------------------------------------------------------------------
int igs_load_gbdt(igs_engine* e, const igs_gbdt_desc* d) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, d != nullptr, "desc is null");
    IGS_REQUIRE(e, d->n_features >= 1 && d->n_features <= IGS_NUM_FEATURES, "n_features must be 1..64");
    IGS_REQUIRE(e, d->post_transform <= IGS_POST_LOGISTIC, "post_transform");
    IGS_REQUIRE(e, d->n_trees <= (1u << 20), "n_trees must be <= 2^20");
    IGS_REQUIRE(e, d->n_trees == 0 || (d->feat && d->thr && d->leaf), "feat/thr/leaf pointer is null");
    IGS_REQUIRE(e, std::isfinite(d->base_score), "base_score must be finite");
    std::lock_guard<std::mutex> lk(e->mu);
    // host-side repack into the device layout (see igs_internal.cuh)
    const size_t bytes = (size_t)d->n_trees * IGS_DEV_TREE_BYTES;
    std::vector<uint8_t> img(bytes ? bytes : 16, 0), img_ws(bytes ? bytes : 16, 0);
    auto ws_feat = [](uint32_t f) { return (uint32_t)igs_ws_row((int)f); };
    for (uint32_t t = 0; t < d->n_trees; ++t) {
        uint8_t* tb = img.data() + (size_t)t * IGS_DEV_TREE_BYTES;
        uint8_t* tw = img_ws.data() + (size_t)t * IGS_DEV_TREE_BYTES;
        const uint8_t* feat = d->feat + (size_t)t * IGS_TREE_NODES;
        const float* thr = d->thr + (size_t)t * IGS_TREE_NODES;
        for (int i = 0; i < IGS_TREE_NODES; ++i)
            if (feat[i] >= d->n_features) {
                igs_set_error(e, "tree %u node %d: feature index %u >= n_features %u", t, i, feat[i], d->n_features);
                return IGS_ERR_INVALID;
            }
        for (int lvl = 0; lvl < IGS_TREE_DEPTH; ++lvl) {
            const int first = (1 << lvl) - 1, count = 1 << lvl;
            for (int j = 0; j < count; ++j) {
                const int i = first + j;
                if ((lvl & 1) == 0) {
                    igs_dev_node nd;
                    nd.feats = ((uint32_t)feat[i] << 7) | ((uint32_t)feat[2 * i + 1] << 13) | ((uint32_t)feat[2 * i + 2] << 19);
                    nd.thr = thr[i];
                    memcpy(tb + igs_dev_level_off(lvl) + j * 8, &nd, 8);
                    nd.feats = (ws_feat(feat[i]) << 7) | (ws_feat(feat[2 * i + 1]) << 13) | (ws_feat(feat[2 * i + 2]) << 19);
                    memcpy(tw + igs_dev_level_off(lvl) + j * 8, &nd, 8);
                } else {
                    memcpy(tb + igs_dev_level_off(lvl) + j * 4, &thr[i], 4);
                    memcpy(tw + igs_dev_level_off(lvl) + j * 4, &thr[i], 4);
                }
            }
        }
        float* leaves = reinterpret_cast<float*>(tb + IGS_DEV_LEAF_OFF);
        for (int i = 0; i < IGS_TREE_LEAVES; ++i) {
            const float lf = d->leaf[(size_t)t * IGS_TREE_LEAVES + i];
            if (!std::isfinite(lf) || fabsf(lf) > 1e30f) {
                igs_set_error(e, "tree %u leaf %d is not finite or exceeds 1e30", t, i);
                return IGS_ERR_INVALID;
            }
            leaves[i] = lf;
        }
        memcpy(tw + IGS_DEV_LEAF_OFF, tb + IGS_DEV_LEAF_OFF, IGS_TREE_LEAVES * sizeof(float));
    }
    IGS_CUDA(e, cudaDeviceSynchronize());
    if (e->d_model) {
        IGS_CUDA(e, cudaFree(e->d_model));
        e->d_model = nullptr;
    }
    if (e->d_model_ws) {
        IGS_CUDA(e, cudaFree(e->d_model_ws));
        e->d_model_ws = nullptr;
    }
    e->has_gbdt = false;
    IGS_CUDA(e, cudaMalloc(&e->d_model, img.size()));
    IGS_CUDA(e, cudaMemcpy(e->d_model, img.data(), img.size(), cudaMemcpyHostToDevice));
    IGS_CUDA(e, cudaMalloc(&e->d_model_ws, img_ws.size()));
    IGS_CUDA(e, cudaMemcpy(e->d_model_ws, img_ws.data(), img_ws.size(), cudaMemcpyHostToDevice));
    e->n_trees = d->n_trees;
    e->n_features = d->n_features;
    e->post_transform = d->post_transform;
    e->base_score = d->base_score;
    e->has_gbdt = true;
    return IGS_OK;
}

int igs_load_mlp(igs_engine* e, const igs_mlp_desc* d) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, d != nullptr, "desc is null");
    std::lock_guard<std::mutex> lk(e->mu);
    return igs_mlp_upload(e, d);
}

int igs_load_players(igs_engine* e, const igs_player_row* rows, uint64_t n) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, n == 0 || rows != nullptr, "rows is null");
    IGS_REQUIRE(e, n < (1ull << 32), "at most 2^32-1 players per shard");
    std::lock_guard<std::mutex> lk(e->mu);
    IGS_CUDA(e, cudaDeviceSynchronize());
    if (e->d_players) {
        IGS_CUDA(e, cudaFree(e->d_players));
        e->d_players = nullptr;
    }
    e->n_players = 0;
    if (n) {
        IGS_CUDA(e, cudaMalloc(&e->d_players, n * sizeof(igs_player_row)));
        IGS_CUDA(e, cudaMemcpy(e->d_players, rows, n * sizeof(igs_player_row), cudaMemcpyHostToDevice));
    }
    e->n_players = n;
    return IGS_OK;
}

int igs_load_events(igs_engine* e, const igs_event_cols* ev, uint64_t n, uint32_t n_accounts) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, ev != nullptr, "event columns are null");
    IGS_REQUIRE(e, n == 0 || (ev->account && ev->ts && ev->amount && ev->tx_type), "an event column is null");
    IGS_REQUIRE(e, n < 0xFFFFFFFFull, "at most 2^32-2 events per shard");
    std::lock_guard<std::mutex> lk(e->mu);
    return igs_build_event_index(e, ev, nullptr, n, n_accounts);
}

int igs_load_events_ids(igs_engine* e, const igs_event_cols* ev, const igs_event_ids* ids, uint64_t n,
                        uint32_t n_accounts) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, ev != nullptr && ids != nullptr, "event / identifier columns are null");
    IGS_REQUIRE(e, n == 0 || (ev->account && ev->ts && ev->amount && ev->tx_type), "an event column is null");
    IGS_REQUIRE(e, n < 0xFFFFFFFFull, "at most 2^32-2 events per shard");
    std::lock_guard<std::mutex> lk(e->mu);
    return igs_build_event_index(e, ev, ids, n, n_accounts);
}

int igs_load_events_device(igs_engine* e, const igs_event_cols* d_ev, const igs_event_ids* d_ids, uint64_t n,
                           uint32_t n_accounts) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, d_ev != nullptr, "event columns are null");
    IGS_REQUIRE(e, n == 0 || (d_ev->account && d_ev->ts && d_ev->amount && d_ev->tx_type), "an event column is null");
    IGS_REQUIRE(e, n < 0xFFFFFFFFull, "at most 2^32-2 events per shard");
    std::lock_guard<std::mutex> lk(e->mu);
    return igs_build_event_index(e, d_ev, d_ids, n, n_accounts);
}

int igs_append_events_ids(igs_engine* e, const igs_event_cols* ev, const igs_event_ids* ids, uint64_t n) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, ev != nullptr && ids != nullptr, "event / identifier columns are null");
    IGS_REQUIRE(e, n == 0 || (ev->account && ev->ts && ev->amount && ev->tx_type), "an event column is null");
    std::lock_guard<std::mutex> lk(e->mu);
    return igs_append_event_columns(e, ev, ids, n);
}

int igs_evict_events(igs_engine* e, uint32_t now, uint64_t* n_evicted) {
    IGS_ENGINE(e);
    std::lock_guard<std::mutex> lk(e->mu);
    return igs_evict_event_history(e, now, n_evicted);
}

int igs_set_late_policy(igs_engine* e, int policy) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, policy == IGS_LATE_REJECT || policy == IGS_LATE_REORDER, "policy must be IGS_LATE_REJECT or IGS_LATE_REORDER");
    std::lock_guard<std::mutex> lk(e->mu);
    e->late_policy = policy;
    return IGS_OK;
}

int igs_set_compaction_threshold(igs_engine* e, uint64_t min_delta_events) {
    IGS_ENGINE(e);
    std::lock_guard<std::mutex> lk(e->mu);
    e->delta_min = min_delta_events;
    return IGS_OK;
}

int igs_prepare_ingest(igs_engine* e, uint64_t max_batch_events) {
    IGS_ENGINE(e);
    std::lock_guard<std::mutex> lk(e->mu);
    return igs_prepare_event_ingest(e, max_batch_events);
}

int igs_get_history_info(igs_engine* e, igs_history_info* out) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, out != nullptr, "out is null");
    std::lock_guard<std::mutex> lk(e->mu);
    memset(out, 0, sizeof(*out));
    if (!e->has_events) return IGS_OK;
    out->n_events = e->n_raw;
    out->late_events = e->late_events;
    out->n_accounts = e->ev.n_accounts;
    out->delta_events = e->n_raw - e->base_n;
    out->has_identifiers = e->has_event_ids ? 1u : 0u;
    return IGS_OK;
}

static int blacklist_update(igs_engine* e, int kind, const uint64_t* ids, uint64_t n, bool replace) {
    IGS_ENGINE(e);
    if (kind != IGS_BL_DEVICE && kind != IGS_BL_FINGERPRINT && kind != IGS_BL_IP) {
        igs_set_error(e, "unknown blacklist type: %d", kind);       // features/redis_store.go:261
        return IGS_ERR_INVALID;
    }
    IGS_REQUIRE(e, n == 0 || ids != nullptr, "ids is null");
    std::lock_guard<std::mutex> lk(e->mu);
    return igs_idset_load(e, &e->bl[kind], ids, n, replace);
}
int igs_blacklist_load(igs_engine* e, int kind, const uint64_t* ids, uint64_t n) {
    return blacklist_update(e, kind, ids, n, true);
}
int igs_blacklist_add(igs_engine* e, int kind, const uint64_t* ids, uint64_t n) {
    return blacklist_update(e, kind, ids, n, false);
}

}  // extern "C"

int igs_append_events(igs_engine* e, const igs_event_cols* ev, uint64_t n) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, ev != nullptr, "event columns are null");
    IGS_REQUIRE(e, n == 0 || (ev->account && ev->ts && ev->amount && ev->tx_type), "an event column is null");
    std::lock_guard<std::mutex> lk(e->mu);
    return igs_append_event_columns(e, ev, nullptr, n);
}

// ------------------------------------------------------------------------------------------------
// chunk pipeline of the host-buffer calls
// ------------------------------------------------------------------------------------------------
// A host-buffer call is PCIe-bound (path A moves 312 B/tx against ~2 ns/tx of kernel time), so the
// batch is cut into chunks that go round-robin over the slot's worker streams: the H2D copy of chunk
// k+1, the kernel of chunk k and the D2H copy of chunk k-1 overlap.  The workers fork from and join
// back into the slot's stream, so ordering against earlier calls on the slot and igs_wait() hold.
constexpr uint64_t kPipeChunkRowsA = 128 * 1024;   // path A: 312 B/tx -> 40 MB per chunk
// Path B: 40 B/tx.  (Chunks large enough for the account-ordered visit of k_score.cu — 256K rows — leave the three
// pipeline streams too few pieces to overlap: measured 7 % slower end to end than 128K-row chunks without it.)
constexpr uint64_t kPipeChunkRowsB = 128 * 1024;

template <typename Body>
static int run_pipelined(igs_engine* e, igs_slot& sl, uint64_t n, uint64_t kPipeChunkRows,
                         Body&& body /* (stream, off, cnt) -> int */) {
    if (n < 2 * kPipeChunkRows) return body(sl.stream, (uint64_t)0, n);
    IGS_CUDA(e, cudaEventRecord(sl.ev_fork, sl.stream));
    for (int k = 0; k < IGS_PIPE_STREAMS; ++k) IGS_CUDA(e, cudaStreamWaitEvent(sl.pipe[k], sl.ev_fork, 0));
    int k = 0, rc = IGS_OK;
    for (uint64_t off = 0; off < n && rc == IGS_OK; off += kPipeChunkRows, k = (k + 1) % IGS_PIPE_STREAMS) {
        const uint64_t cnt = std::min(kPipeChunkRows, n - off);
        rc = body(sl.pipe[k], off, cnt);
    }
    // Join the workers into the slot's stream whether or not a chunk failed: chunks already queued keep copying
    // into the caller's buffers, and igs_wait / the synchronous wrappers must cover them.
    for (int j = 0; j < IGS_PIPE_STREAMS; ++j) {
        const cudaError_t a = cudaEventRecord(sl.ev_join[j], sl.pipe[j]);
        const cudaError_t b = a == cudaSuccess ? cudaStreamWaitEvent(sl.stream, sl.ev_join[j], 0) : a;
        if (b != cudaSuccess) {
            cudaStreamSynchronize(sl.pipe[j]);          // could not order it: wait for it here
            if (rc == IGS_OK) {
                igs_set_error(e, "joining the chunk pipeline failed: %s", cudaGetErrorString(b));
                rc = IGS_ERR_CUDA;
            }
        }
    }
    return rc;
}

// Synchronous wrappers: wait for the slot's stream even when queueing failed half way, so that nothing is still
// writing into the caller's buffers when the call returns; the first error wins.
static int finish_sync(igs_engine* e, int slot, int rc) {
    const cudaError_t st = cudaStreamSynchronize(e->slot[slot].stream);
    if (rc != IGS_OK) return rc;
    if (st != cudaSuccess) {
        igs_set_error(e, "cudaStreamSynchronize failed: %s", cudaGetErrorString(st));
        return IGS_ERR_CUDA;
    }
    return IGS_OK;
}

extern "C" {

// ------------------------------------------------------------------------------------------------
// path A
// ------------------------------------------------------------------------------------------------
// Host-buffer scoring on one slot: the batch is cut into chunks; each chunk's H2D copy, kernel and
// D2H copy are queued on the slot's stream, so chunk k+1 uploads while chunk k computes only when
// the caller uses several slots.  (One stream keeps result order trivially correct.)
static int score_features_host(igs_engine* e, int s, const igs_config* cfg, const float* features,
                               const igs_rule_row* rows, uint64_t n, igs_result* out, uint8_t* leaf_idx) {
    IGS_REQUIRE(e, cfg != nullptr, "cfg is null");
    IGS_REQUIRE(e, n == 0 || (rows && out), "rows / out is null");
    IGS_REQUIRE(e, n == 0 || cfg->model_kind == IGS_MODEL_NONE || features, "features is null");
    if (n == 0) return IGS_OK;
    igs_slot& sl = e->slot[s];
    const bool need_feat = cfg->model_kind != IGS_MODEL_NONE;
    const bool leaves = leaf_idx != nullptr && cfg->model_kind == IGS_MODEL_GBDT;
    const uint32_t nt = e->n_trees;
    IGS_TRY(igs_reserve(e, &sl.in0, need_feat ? n * IGS_NUM_FEATURES * sizeof(float) : 0));
    IGS_TRY(igs_reserve(e, &sl.in1, n * sizeof(igs_rule_row)));
    IGS_TRY(igs_reserve(e, &sl.out0, n * sizeof(igs_result)));
    if (leaves) IGS_TRY(igs_reserve(e, &sl.aux0, n * (size_t)nt));
    float* d_feat = (float*)sl.in0.p;
    igs_rule_row* d_rows = (igs_rule_row*)sl.in1.p;
    igs_result* d_out = (igs_result*)sl.out0.p;
    uint8_t* d_leaf = leaves ? (uint8_t*)sl.aux0.p : nullptr;
    return run_pipelined(e, sl, n, kPipeChunkRowsA, [&](cudaStream_t st, uint64_t off, uint64_t cnt) -> int {
        if (need_feat)
            IGS_CUDA(e, cudaMemcpyAsync(d_feat + off * IGS_NUM_FEATURES, features + off * IGS_NUM_FEATURES,
                                        cnt * IGS_NUM_FEATURES * sizeof(float), cudaMemcpyHostToDevice, st));
        IGS_CUDA(e, cudaMemcpyAsync(d_rows + off, rows + off, cnt * sizeof(igs_rule_row), cudaMemcpyHostToDevice, st));
        IGS_TRY(igs_launch_score_features(e, cfg, d_feat + off * IGS_NUM_FEATURES, d_rows + off, cnt, d_out + off,
                                          leaves ? d_leaf + off * nt : nullptr, st));
        IGS_CUDA(e, cudaMemcpyAsync(out + off, d_out + off, cnt * sizeof(igs_result), cudaMemcpyDeviceToHost, st));
        if (leaves)
            IGS_CUDA(e, cudaMemcpyAsync(leaf_idx + off * nt, d_leaf + off * nt, cnt * (size_t)nt,
                                        cudaMemcpyDeviceToHost, st));
        return IGS_OK;
    });
}

int igs_score_features(igs_engine* e, const igs_config* cfg, const float* features,
                       const igs_rule_row* rows, uint64_t n, igs_result* out, uint8_t* leaf_idx) {
    IGS_ENGINE(e);
    return finish_sync(e, 0, score_features_host(e, 0, cfg, features, rows, n, out, leaf_idx));
}

int igs_score_features_async(igs_engine* e, int slot, const igs_config* cfg, const float* features,
                             const igs_rule_row* rows, uint64_t n, igs_result* out) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, slot >= 0 && slot < IGS_NUM_SLOTS, "slot out of range");
    return score_features_host(e, slot, cfg, features, rows, n, out, nullptr);
}

int igs_wait(igs_engine* e, int slot) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, slot >= 0 && slot < IGS_NUM_SLOTS, "slot out of range");
    IGS_CUDA(e, cudaStreamSynchronize(e->slot[slot].stream));
    return IGS_OK;
}

int igs_score_features_device(igs_engine* e, const igs_config* cfg, const float* d_features,
                              const igs_rule_row* d_rows, uint64_t n, igs_result* d_out,
                              uint8_t* d_leaf_idx, void* stream) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, cfg != nullptr, "cfg is null");
    return igs_launch_score_features(e, cfg, d_features, d_rows, n, d_out, d_leaf_idx, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// velocity
// ------------------------------------------------------------------------------------------------
int igs_get_velocity(igs_engine* e, const igs_query* q, uint64_t n, igs_velocity* out) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, n == 0 || (q && out), "q / out is null");
    if (n == 0) return IGS_OK;
    igs_slot& sl = e->slot[0];
    IGS_TRY(igs_reserve(e, &sl.in0, n * sizeof(igs_query)));
    IGS_TRY(igs_reserve(e, &sl.out0, n * sizeof(igs_velocity)));
    IGS_CUDA(e, cudaMemcpyAsync(sl.in0.p, q, n * sizeof(igs_query), cudaMemcpyHostToDevice, sl.stream));
    IGS_TRY(igs_launch_velocity(e, (const igs_query*)sl.in0.p, n, (igs_velocity*)sl.out0.p, sl.stream));
    IGS_CUDA(e, cudaMemcpyAsync(out, sl.out0.p, n * sizeof(igs_velocity), cudaMemcpyDeviceToHost, sl.stream));
    IGS_CUDA(e, cudaStreamSynchronize(sl.stream));
    return IGS_OK;
}

int igs_get_velocity_device(igs_engine* e, const igs_query* d_q, uint64_t n, igs_velocity* d_out,
                            void* stream) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, n == 0 || (d_q && d_out), "q / out is null");
    return igs_launch_velocity(e, d_q, n, d_out, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// path B
// ------------------------------------------------------------------------------------------------
// pinned host memory that the device can address through the same pointer (cudaHostAlloc / igs_alloc_pinned under UVA)
static bool host_mapped(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost && a.devicePointer == p;
}

// Large batches in pinned host buffers: ONE kernel launch over the whole batch, started at once; the records follow
// it into device memory in IGS_ARRIVAL_CHUNKS copies on a worker stream, each followed by a 4-byte copy that bumps the
// slot's arrival counter, and the kernel's producer warps wait for the chunk they are about to read (igs_tx_arrival).
// The copy engine runs ahead of the kernel after the first chunks (32 MB at ~55 GB/s against 1.7 ms of scoring), so
// the producers read HBM, not PCIe; the 8-byte results are written straight to the caller's pinned buffer.
static int score_batch_streamed(igs_engine* e, igs_slot& sl, const igs_config* cfg, const igs_tx* tx,
                                const igs_tx_ids* ids, uint64_t n, igs_result* out) {
    IGS_TRY(igs_reserve(e, &sl.in0, n * sizeof(igs_tx)));
    igs_tx* d_tx = (igs_tx*)sl.in0.p;
    // chunks of whole 512-row tiles
    const uint64_t rows = ((n + IGS_ARRIVAL_CHUNKS - 1) / IGS_ARRIVAL_CHUNKS + 511) / 512 * 512;
    IGS_REQUIRE(e, rows <= 0xFFFFFFFFull, "batch too large");
    cudaStream_t cp = sl.pipe[0];
    IGS_CUDA(e, cudaMemsetAsync(sl.d_ready, 0, sizeof(uint32_t), sl.stream));
    IGS_CUDA(e, cudaEventRecord(sl.ev_fork, sl.stream));
    IGS_CUDA(e, cudaStreamWaitEvent(cp, sl.ev_fork, 0));
    const igs_tx_arrival arrival = {sl.d_ready, (uint32_t)rows};
    int rc = igs_launch_score_batch(e, cfg, d_tx, ids, n, out, nullptr, nullptr, sl.stream, &arrival);
    if (rc != IGS_OK) return rc;                       // nothing was launched
    // From here on the kernel is waiting for the copies: every chunk must be queued, whatever happens.
    cudaError_t st = cudaSuccess;
    int c = 0;
    for (uint64_t off = 0; off < n && st == cudaSuccess; off += rows, ++c) {
        const uint64_t cnt = std::min(rows, n - off);
        st = cudaMemcpyAsync(d_tx + off, tx + off, cnt * sizeof(igs_tx), cudaMemcpyHostToDevice, cp);
        if (st == cudaSuccess)
            st = cudaMemcpyAsync(sl.d_ready, e->h_arrival_vals + c, sizeof(uint32_t), cudaMemcpyHostToDevice, cp);
    }
    if (st != cudaSuccess) {
        // release the kernel (it scores whatever is in the staging buffer; the call fails anyway)
        cudaMemcpyAsync(sl.d_ready, e->h_arrival_vals + IGS_ARRIVAL_CHUNKS - 1, sizeof(uint32_t), cudaMemcpyHostToDevice, cp);
        igs_set_error(e, "queueing the batch copies failed: %s", cudaGetErrorString(st));
        rc = IGS_ERR_CUDA;
    }
    const cudaError_t a = cudaEventRecord(sl.ev_join[0], cp);
    const cudaError_t b = a == cudaSuccess ? cudaStreamWaitEvent(sl.stream, sl.ev_join[0], 0) : a;
    if (b != cudaSuccess) cudaStreamSynchronize(cp);
    return rc;
}

static int score_batch_host(igs_engine* e, int s, const igs_config* cfg, const igs_tx* tx, const igs_tx_ids* ids,
                            uint64_t n, igs_result* out, float* features_out, igs_velocity* velocity_out) {
    IGS_REQUIRE(e, cfg != nullptr, "cfg is null");
    IGS_REQUIRE(e, n == 0 || (tx && out), "tx / out is null");
    if (n == 0) return IGS_OK;
    igs_slot& sl = e->slot[s];
    // Large batches in pinned (mapped) host buffers are not cut into chunk launches: one kernel over the whole batch
    // with the records streaming in behind its start (score_batch_streamed); results go straight to the host buffer.
    if (!features_out && !velocity_out && igs_fused_direct_eligible(e, n) && host_mapped(tx) && host_mapped(out) &&
        (!ids || host_mapped(ids))) {
        static const int direct = getenv("IGS_SCORE_DIRECT") ? atoi(getenv("IGS_SCORE_DIRECT")) : 1;   // tuning knob
        // (a kernel that waits for copies queued after its launch cannot run under CUDA_LAUNCH_BLOCKING=1)
        static const bool blocking = getenv("CUDA_LAUNCH_BLOCKING") && atoi(getenv("CUDA_LAUNCH_BLOCKING")) != 0;
        if (direct == 2 || blocking)   // the producers read the records themselves, over PCIe
            return igs_launch_score_batch(e, cfg, tx, ids, n, out, nullptr, nullptr, sl.stream, nullptr, true);
        return score_batch_streamed(e, sl, cfg, tx, ids, n, out);
    }
    IGS_TRY(igs_reserve(e, &sl.in0, n * sizeof(igs_tx)));
    if (ids) IGS_TRY(igs_reserve(e, &sl.in1, n * sizeof(igs_tx_ids)));
    igs_tx_ids* d_ids = ids ? (igs_tx_ids*)sl.in1.p : nullptr;
    IGS_TRY(igs_reserve(e, &sl.out0, n * sizeof(igs_result)));
    if (features_out) IGS_TRY(igs_reserve(e, &sl.aux0, n * IGS_NUM_FEATURES * sizeof(float)));
    if (velocity_out) IGS_TRY(igs_reserve(e, &sl.aux1, n * sizeof(igs_velocity)));
    igs_tx* d_tx = (igs_tx*)sl.in0.p;
    igs_result* d_out = (igs_result*)sl.out0.p;
    float* d_feat = features_out ? (float*)sl.aux0.p : nullptr;
    igs_velocity* d_vel = velocity_out ? (igs_velocity*)sl.aux1.p : nullptr;
    // Chunks large enough for the warp-specialised kernel (>= 4 tiles of 512 tx per SM, k_score.cu) and equal in size:
    // 1M tx on 148 SMs = three chunks of 333 334 rows.
    const uint64_t ws_rows = 4ull * (uint64_t)e->sm_count * 512ull;
    const uint64_t pieces = std::max<uint64_t>(1, n / (ws_rows + ws_rows / 16));
    const uint64_t chunk_rows = std::max<uint64_t>(kPipeChunkRowsB, (n + pieces - 1) / pieces);
    return run_pipelined(e, sl, n, chunk_rows, [&](cudaStream_t st, uint64_t off, uint64_t cnt) -> int {
        IGS_CUDA(e, cudaMemcpyAsync(d_tx + off, tx + off, cnt * sizeof(igs_tx), cudaMemcpyHostToDevice, st));
        if (ids)
            IGS_CUDA(e, cudaMemcpyAsync(d_ids + off, ids + off, cnt * sizeof(igs_tx_ids), cudaMemcpyHostToDevice, st));
        IGS_TRY(igs_launch_score_batch(e, cfg, d_tx + off, ids ? d_ids + off : nullptr, cnt, d_out + off,
                                       d_feat ? d_feat + off * IGS_NUM_FEATURES : nullptr,
                                       d_vel ? d_vel + off : nullptr, st));
        IGS_CUDA(e, cudaMemcpyAsync(out + off, d_out + off, cnt * sizeof(igs_result), cudaMemcpyDeviceToHost, st));
        if (features_out)
            IGS_CUDA(e, cudaMemcpyAsync(features_out + off * IGS_NUM_FEATURES, d_feat + off * IGS_NUM_FEATURES,
                                        cnt * IGS_NUM_FEATURES * sizeof(float), cudaMemcpyDeviceToHost, st));
        if (velocity_out)
            IGS_CUDA(e, cudaMemcpyAsync(velocity_out + off, d_vel + off, cnt * sizeof(igs_velocity),
                                        cudaMemcpyDeviceToHost, st));
        return IGS_OK;
    });
}

int igs_score_batch(igs_engine* e, const igs_config* cfg, const igs_tx* tx, uint64_t n, igs_result* out,
                    float* features_out, igs_velocity* velocity_out) {
    IGS_ENGINE(e);
    return finish_sync(e, 0, score_batch_host(e, 0, cfg, tx, nullptr, n, out, features_out, velocity_out));
}

int igs_score_batch_async(igs_engine* e, int slot, const igs_config* cfg, const igs_tx* tx, uint64_t n,
                          igs_result* out) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, slot >= 0 && slot < IGS_NUM_SLOTS, "slot out of range");
    return score_batch_host(e, slot, cfg, tx, nullptr, n, out, nullptr, nullptr);
}

int igs_score_batch_device(igs_engine* e, const igs_config* cfg, const igs_tx* d_tx, uint64_t n,
                           igs_result* d_out, float* d_features_out, igs_velocity* d_velocity_out,
                           void* stream) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, cfg != nullptr, "cfg is null");
    IGS_REQUIRE(e, n == 0 || (d_tx && d_out), "tx / out is null");
    return igs_launch_score_batch(e, cfg, d_tx, nullptr, n, d_out, d_features_out, d_velocity_out,
                                  (cudaStream_t)stream);
}

int igs_score_batch_ids(igs_engine* e, const igs_config* cfg, const igs_tx* tx, const igs_tx_ids* ids, uint64_t n,
                        igs_result* out, float* features_out, igs_velocity* velocity_out) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, n == 0 || ids != nullptr, "ids is null");
    return finish_sync(e, 0, score_batch_host(e, 0, cfg, tx, ids, n, out, features_out, velocity_out));
}

int igs_score_batch_ids_device(igs_engine* e, const igs_config* cfg, const igs_tx* d_tx, const igs_tx_ids* d_ids,
                               uint64_t n, igs_result* d_out, float* d_features_out,
                               igs_velocity* d_velocity_out, void* stream) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, cfg != nullptr, "cfg is null");
    IGS_REQUIRE(e, n == 0 || (d_tx && d_ids && d_out), "tx / ids / out is null");
    return igs_launch_score_batch(e, cfg, d_tx, d_ids, n, d_out, d_features_out, d_velocity_out,
                                  (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// identifiers: uniques and blacklist checks
// ------------------------------------------------------------------------------------------------
int igs_get_uniques(igs_engine* e, const igs_query* q, uint64_t n, igs_uniques* out) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, n == 0 || (q && out), "q / out is null");
    if (n == 0) return IGS_OK;
    igs_slot& sl = e->slot[0];
    IGS_TRY(igs_reserve(e, &sl.in0, n * sizeof(igs_query)));
    IGS_TRY(igs_reserve(e, &sl.out0, n * sizeof(igs_uniques)));
    IGS_CUDA(e, cudaMemcpyAsync(sl.in0.p, q, n * sizeof(igs_query), cudaMemcpyHostToDevice, sl.stream));
    IGS_TRY(igs_launch_uniques(e, (const igs_query*)sl.in0.p, n, (igs_uniques*)sl.out0.p, sl.stream));
    IGS_CUDA(e, cudaMemcpyAsync(out, sl.out0.p, n * sizeof(igs_uniques), cudaMemcpyDeviceToHost, sl.stream));
    IGS_CUDA(e, cudaStreamSynchronize(sl.stream));
    return IGS_OK;
}

int igs_get_uniques_device(igs_engine* e, const igs_query* d_q, uint64_t n, igs_uniques* d_out, void* stream) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, n == 0 || (d_q && d_out), "q / out is null");
    return igs_launch_uniques(e, d_q, n, d_out, (cudaStream_t)stream);
}

int igs_blacklist_check(igs_engine* e, const igs_tx_ids* ids, uint64_t n, uint8_t* out) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, n == 0 || (ids && out), "ids / out is null");
    if (n == 0) return IGS_OK;
    igs_slot& sl = e->slot[0];
    IGS_TRY(igs_reserve(e, &sl.in1, n * sizeof(igs_tx_ids)));
    IGS_TRY(igs_reserve(e, &sl.out0, n));
    IGS_CUDA(e, cudaMemcpyAsync(sl.in1.p, ids, n * sizeof(igs_tx_ids), cudaMemcpyHostToDevice, sl.stream));
    IGS_TRY(igs_launch_blacklist_check(e, (const igs_tx_ids*)sl.in1.p, n, (uint8_t*)sl.out0.p, sl.stream));
    IGS_CUDA(e, cudaMemcpyAsync(out, sl.out0.p, n, cudaMemcpyDeviceToHost, sl.stream));
    IGS_CUDA(e, cudaStreamSynchronize(sl.stream));
    return IGS_OK;
}

int igs_blacklist_check_device(igs_engine* e, const igs_tx_ids* d_ids, uint64_t n, uint8_t* d_out, void* stream) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, n == 0 || (d_ids && d_out), "ids / out is null");
    return igs_launch_blacklist_check(e, d_ids, n, d_out, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// shard partition
// ------------------------------------------------------------------------------------------------
int igs_partition_tx_device(igs_engine* e, const igs_tx* d_tx, uint64_t n, uint32_t n_shards,
                            igs_tx* d_tx_out, uint32_t* d_perm, uint32_t* d_offsets, void* stream) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, n_shards >= 1 && n_shards <= 256, "n_shards must be 1..256");
    IGS_REQUIRE(e, n < (1ull << 32), "at most 2^32-1 records per call");
    IGS_REQUIRE(e, d_offsets != nullptr, "offsets is null");
    IGS_REQUIRE(e, n == 0 || (d_tx && d_tx_out), "tx / tx_out is null");
    {
        std::lock_guard<std::mutex> lk(e->mu);
        IGS_TRY(igs_reserve(e, &e->part_scratch, igs_partition_scratch_words(n, n_shards) * sizeof(uint32_t)));
    }
    return igs_partition_run(e, d_tx, n, n_shards, (uint32_t*)e->part_scratch.p, d_tx_out, d_perm, d_offsets,
                             (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// LTV
// ------------------------------------------------------------------------------------------------
int igs_ltv_batch(igs_engine* e, const uint16_t* x, const igs_ltv_row* rows, uint64_t n,
                  igs_ltv_result* out, float* ltv_raw) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, n == 0 || (x && rows && out), "x / rows / out is null");
    if (n == 0) return IGS_OK;
    igs_slot& sl = e->slot[0];
    IGS_TRY(igs_reserve(e, &sl.in0, n * IGS_MLP_IN * sizeof(uint16_t)));
    IGS_TRY(igs_reserve(e, &sl.in1, n * sizeof(igs_ltv_row)));
    IGS_TRY(igs_reserve(e, &sl.out0, n * sizeof(igs_ltv_result)));
    IGS_TRY(igs_reserve(e, &sl.aux0, n * sizeof(float)));
    IGS_CUDA(e, cudaMemcpyAsync(sl.in0.p, x, n * IGS_MLP_IN * sizeof(uint16_t), cudaMemcpyHostToDevice, sl.stream));
    IGS_CUDA(e, cudaMemcpyAsync(sl.in1.p, rows, n * sizeof(igs_ltv_row), cudaMemcpyHostToDevice, sl.stream));
    IGS_TRY(igs_launch_ltv(e, (const uint16_t*)sl.in0.p, (const igs_ltv_row*)sl.in1.p, n,
                           (igs_ltv_result*)sl.out0.p, (float*)sl.aux0.p, sl.stream));
    IGS_CUDA(e, cudaMemcpyAsync(out, sl.out0.p, n * sizeof(igs_ltv_result), cudaMemcpyDeviceToHost, sl.stream));
    if (ltv_raw)
        IGS_CUDA(e, cudaMemcpyAsync(ltv_raw, sl.aux0.p, n * sizeof(float), cudaMemcpyDeviceToHost, sl.stream));
    IGS_CUDA(e, cudaStreamSynchronize(sl.stream));
    return IGS_OK;
}

int igs_ltv_batch_device(igs_engine* e, const uint16_t* d_x, const igs_ltv_row* d_rows, uint64_t n,
                         igs_ltv_result* d_out, float* d_ltv_raw, void* stream) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, n == 0 || (d_x && d_rows && d_out), "x / rows / out is null");
    return igs_launch_ltv(e, d_x, d_rows, n, d_out, d_ltv_raw, (cudaStream_t)stream);
}

int igs_ltv_hidden_device(igs_engine* e, const uint16_t* d_x, uint64_t n, int layer, uint16_t* d_hidden,
                          void* stream) {
    IGS_ENGINE(e);
    IGS_REQUIRE(e, n == 0 || (d_x && d_hidden), "x / hidden is null");
    return igs_launch_ltv_hidden(e, d_x, n, layer, d_hidden, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// host-side helpers
// ------------------------------------------------------------------------------------------------
void igs_default_config(igs_config* c) {
    if (!c) return;
    memset(c, 0, sizeof(*c));
    c->block_threshold = 80;       // scoring/engine.go:217
    c->review_threshold = 50;      // :218
    c->max_tx_per_minute = 10;     // :219
    c->max_tx_per_hour = 100;      // :220
    c->new_account_days = 7;       // :221
    c->large_deposit_amount = 100000;  // :222
    c->max_devices_per_day = 3;    // :223
    c->max_ips_per_day = 5;        // :224
    c->ml_weight = 0.6;            // :225
    c->rule_weight = 0.4;          // :226
    c->model_kind = IGS_MODEL_MOCK;
    static const int32_t w[IGS_NUM_REASONS] = {20, 30, 25, 15, 20, 15, 50, 25, 20, 30};  // :246-257
    memcpy(c->reason_weight, w, sizeof(w));
}

int32_t igs_rule_score_from_flags(const igs_config* cfg, uint32_t flags) {
    if (!cfg) return 0;
    int32_t total = 0;
    for (int b = 0; b < IGS_NUM_REASONS; ++b)
        if ((flags >> b) & 1u)
            if ((1u << b) != IGS_R_ML_HIGH_RISK) total += cfg->reason_weight[b];
    return total > 100 ? 100 : total;
}

int igs_reason_codes(uint32_t flags, const char* out[IGS_NUM_REASONS]) {
    // emission order of applyRules (scoring/engine.go:424-475), then the ML reason (:285-287)
    static const struct { uint32_t bit; const char* name; } order[] = {
        {IGS_R_HIGH_VELOCITY, "HIGH_VELOCITY"},
        {IGS_R_NEW_ACCOUNT_LARGE_TX, "NEW_ACCOUNT_LARGE_TX"},
        {IGS_R_MULTIPLE_DEVICES, "MULTIPLE_DEVICES"},
        {IGS_R_IP_COUNTRY_MISMATCH, "IP_COUNTRY_MISMATCH"},
        {IGS_R_VPN_DETECTED, "VPN_DETECTED"},
        {IGS_R_RAPID_DEPOSIT_WITHDRAW, "RAPID_DEPOSIT_WITHDRAW"},
        {IGS_R_BONUS_ABUSE, "BONUS_ABUSE"},
        {IGS_R_KNOWN_FRAUDSTER, "KNOWN_FRAUDSTER"},
        {IGS_R_SUSPICIOUS_PATTERN, "SUSPICIOUS_PATTERN"},
        {IGS_R_ML_HIGH_RISK, "ML_HIGH_RISK"},
    };
    int k = 0;
    for (const auto& o : order)
        if (flags & o.bit) out[k++] = o.name;
    return k;
}

const char* igs_action_name(int action) {
    switch (action) {
    case IGS_ACTION_APPROVE: return "approve";
    case IGS_ACTION_REVIEW: return "review";
    case IGS_ACTION_BLOCK: return "block";
    default: return "";
    }
}

const char* igs_ltv_action_name(int idx) {
    static const char* names[] = {
        "NO_ACTION", "SEND_WINBACK_BONUS", "SEND_ENGAGEMENT_EMAIL", "VIP_MANAGER_CALL",
        "EXCLUSIVE_EVENT_INVITE", "ASSIGN_VIP_MANAGER", "RETENTION_BONUS", "LOYALTY_REWARD",
        "SUGGEST_BONUS", "RECOMMEND_NEW_GAMES", "STANDARD_PROMOTION", "ONBOARDING_GUIDE",
        "SMALL_DEPOSIT_BONUS"};
    return idx >= 0 && idx < (int)(sizeof(names) / sizeof(names[0])) ? names[idx] : "";
}

const char* igs_segment_name(int seg) {
    switch (seg) {
    case IGS_SEG_VIP: return "vip";
    case IGS_SEG_HIGH: return "high";
    case IGS_SEG_MEDIUM: return "medium";
    case IGS_SEG_LOW: return "low";
    case IGS_SEG_CHURNING: return "churning";
    default: return "";
    }
}

static uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

uint32_t igs_shard_of(uint64_t lo, uint64_t hi, uint32_t n_shards) {
    return n_shards ? (uint32_t)(splitmix64(lo ^ hi) % n_shards) : 0;
}

unsigned long long igs_kernel_launches(void) { return __atomic_load_n(&g_igs_launches, __ATOMIC_RELAXED); }

const char* igs_version(void) { return "igs 0.1 sm_100a"; }

}  // extern "C"
