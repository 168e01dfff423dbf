/*
 * igs.h — C ABI of the B200 batched fraud-scoring engine ("igs" = iGaming Scoring).
 *
 * This is the drop-in boundary for the scoring path of the reference's Risk & Prediction
 * service (services/risk).  Every entry point names the reference interface it replaces
 * (paths relative to the reference root).  Plain C: pointers, sizes, PODs.  No C++
 * exceptions cross this boundary; every call returns IGS_OK or a negative status and
 * leaves a message retrievable through igs_last_error().
 *
 * Ownership: the caller owns every host buffer (use igs_alloc_pinned() for anything that is
 * passed to a *host-buffer* entry point on the hot path — pageable memory works but is slower).
 * The engine owns device memory, streams and staging.  Inputs are never mutated
 * (the reference mutates the caller's vector in Normalize(), ml/onnx_model.go:213 — we do not).
 *
 * Threading: one engine per device.  Calls on different *slots* of one engine may be issued
 * from different host threads concurrently; calls on the same slot must be serialised by the
 * caller (the synchronous host-buffer calls all use slot 0).  The calls that replace device state —
 * igs_load_*, igs_append_events*, igs_evict_events, igs_blacklist_load / _add, igs_links_* — are
 * like ONNXModel.Close under its write lock (ml/onnx_model.go:349): they must not run while a
 * scoring or query call on the same engine is in flight — with one exception: igs_append_events
 * (history without identifiers) may run concurrently with scoring and query calls, see there.  Three *_device calls work through per-engine
 * scratch, not per-stream scratch: igs_partition_tx_device (its counters; grown with cudaFree +
 * cudaMalloc when a batch needs more than any before), igs_ltv_batch_device and igs_ltv_hidden_device
 * (two 512 MB activation buffers shared by every call).  On one engine, order such calls with respect to
 * each other — one stream, or events between streams; two of them overlapping on the GPU overwrite each
 * other's intermediate data and return wrong results silently.  The scoring *_device calls have no such
 * scratch (the account-ordered visit of large fused batches allocates stream-ordered) and may overlap.  igs_last_error returns the engine's single message buffer; read it on the thread
 * that saw the failing status.  Configuration (weights / thresholds) is passed by value with
 * every call, so there is no shared mutable config (the reference has a latent race there:
 * scoring/engine.go:304-306 vs :498-504).
 *
 * There is NO CPU fallback behind this ABI: without a CUDA device igs_engine_create() fails.
 */
#ifndef IGS_H_
#define IGS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------ */
/* status codes                                                                               */
/* ------------------------------------------------------------------------------------------ */
#define IGS_OK            0
#define IGS_ERR_INVALID  (-1) /* bad argument (null pointer, size out of range, bad enum)      */
#define IGS_ERR_CUDA     (-2) /* a CUDA runtime call failed; igs_last_error() has the text     */
#define IGS_ERR_STATE    (-3) /* call needs state that was not loaded (model, players, events) */
#define IGS_ERR_NOMEM    (-4) /* host or device allocation failed                              */

/* ------------------------------------------------------------------------------------------ */
/* fixed shapes of the path                                                                   */
/* ------------------------------------------------------------------------------------------ */
#define IGS_NUM_FEATURES  64   /* model input width (north-star); idx 0..29 = reference order   */
#define IGS_REF_FEATURES  30   /* ml.FeatureVector, ml/onnx_model.go:86-130 / ToSlice :133-166  */
#define IGS_TREE_DEPTH     8
#define IGS_TREE_NODES   255   /* internal nodes of one complete depth-8 tree, heap order       */
#define IGS_TREE_LEAVES  256
#define IGS_NUM_SLOTS      4   /* independent stream + staging sets per engine                  */

/* Reason bits.  Bit order = ReasonCode const order, scoring/engine.go:17-28.                    */
#define IGS_R_HIGH_VELOCITY           (1u << 0)
#define IGS_R_NEW_ACCOUNT_LARGE_TX    (1u << 1)
#define IGS_R_IP_COUNTRY_MISMATCH     (1u << 2)
#define IGS_R_MULTIPLE_DEVICES        (1u << 3)
#define IGS_R_SUSPICIOUS_PATTERN      (1u << 4) /* declared by the reference, never emitted     */
#define IGS_R_VPN_DETECTED            (1u << 5)
#define IGS_R_KNOWN_FRAUDSTER         (1u << 6)
#define IGS_R_RAPID_DEPOSIT_WITHDRAW  (1u << 7)
#define IGS_R_BONUS_ABUSE             (1u << 8)
#define IGS_R_ML_HIGH_RISK            (1u << 9)
#define IGS_NUM_REASONS 10

/* Action, numbered as proto/risk/v1/risk.proto:84-89.                                          */
#define IGS_ACTION_APPROVE 1
#define IGS_ACTION_REVIEW  2
#define IGS_ACTION_BLOCK   3

/* Transaction type (ScoreRequest.TxType string, scoring/engine.go:44).                         */
#define IGS_TX_BET      0
#define IGS_TX_DEPOSIT  1
#define IGS_TX_WITHDRAW 2
#define IGS_TX_OTHER    3

/* Per-row bits.                                                                               */
#define IGS_BIT_VPN         (1u << 0) /* IPInfo.IsVPN   scoring/engine.go:403                    */
#define IGS_BIT_PROXY       (1u << 1) /* IPInfo.IsProxy :404                                     */
#define IGS_BIT_TOR         (1u << 2) /* IPInfo.IsTor   :405                                     */
#define IGS_BIT_BONUS_ONLY  (1u << 3) /* FeatureVector.BonusOnlyPlayer :384-386 (feature rows)   */
#define IGS_BIT_BLACKLISTED (1u << 4) /* verdict of Blacklist.Check :470 (host-side set lookup)  */
#define IGS_BIT_VALID       (1u << 5) /* 0 = the host could not stage this row's model input:
                                         the model "fails" and ml = 0.5, scoring/engine.go:279-282 */

/* Which model the MLModel slot holds (scoring/engine.go:153-155,276-288).                      */
#define IGS_MODEL_NONE 0 /* model == nil : ml = 0, no ML_HIGH_RISK                              */
#define IGS_MODEL_MOCK 1 /* ml/onnx_model.go:258-308 mockPredict (the only runnable reference)  */
#define IGS_MODEL_GBDT 2 /* tree ensemble loaded with igs_load_gbdt (stands in for fraud_model.onnx) */

/* Output transform of the ensemble margin (ONNX TreeEnsemble post_transform).                 */
#define IGS_POST_NONE     0 /* ml = clamp(margin, 0, 1)                                          */
#define IGS_POST_LOGISTIC 1 /* ml = clamp(igs_sigmoid(margin), 0, 1)                             */

typedef struct igs_engine igs_engine;

/* ------------------------------------------------------------------------------------------ */
/* configuration — mirrors ScoringConfig (scoring/engine.go:196-228) + ruleWeights (:246-257)    */
/* ------------------------------------------------------------------------------------------ */
typedef struct igs_config {
    int32_t block_threshold;       /* >= this score = block   (default 80)  */
    int32_t review_threshold;      /* >= this score = review  (default 50)  */
    int32_t max_tx_per_minute;     /* R1: c1m > this          (default 10)  */
    int32_t max_tx_per_hour;       /* configured, unused by the reference (default 100) */
    int32_t new_account_days;      /* R2: age < this          (default 7)   */
    int32_t max_devices_per_day;   /* R3: devices > this      (default 3)   */
    int32_t max_ips_per_day;       /* R4: ips > this          (default 5)   */
    int32_t model_kind;            /* IGS_MODEL_*                           */
    int64_t large_deposit_amount;  /* R2: amount > this, cents (default 100000) */
    double  ml_weight;             /* default 0.6 */
    double  rule_weight;           /* default 0.4 */
    int32_t reason_weight[IGS_NUM_REASONS]; /* indexed by reason bit number; defaults 20,30,25,15,20,15,50,25,20,30 */
    int32_t reserved[2];
} igs_config;

/* Fills *cfg with DefaultConfig() (scoring/engine.go:215-228) + the constructor's rule weights
 * (:246-257) and model_kind = IGS_MODEL_MOCK (what the reference runs when no .onnx exists,
 * ml/onnx_model.go:50-60).  Pure host code, needs no GPU.                                       */
void igs_default_config(igs_config* cfg);

/* ------------------------------------------------------------------------------------------ */
/* row layouts                                                                                */
/* ------------------------------------------------------------------------------------------ */

/* Rule sideband of one pre-assembled feature row: the exact integers applyRules needs
 * (scoring/engine.go:420-483) — the normalised floats of the model input cannot reproduce
 * the integer comparisons.  48 bytes.                                                          */
typedef struct igs_rule_row {
    int64_t  amount;             /* ScoreRequest.Amount, cents                  engine.go:43   */
    int64_t  total_deposits;     /* FeatureVector.TotalDeposits                 engine.go:83   */
    int64_t  total_withdrawals;  /* FeatureVector.TotalWithdrawals              engine.go:84   */
    uint32_t tx_count_1m;        /* FeatureVector.TxCount1Min                   engine.go:69   */
    int32_t  account_age_days;   /* FeatureVector.AccountAge                    engine.go:82   */
    uint32_t time_since_last_tx; /* FeatureVector.TimeSinceLastTx, seconds      engine.go:90   */
    uint32_t deposit_count;      /* FeatureVector.DepositCount                  engine.go:86   */
    uint16_t unique_devices_24h; /* FeatureVector.UniqueDevices24h              engine.go:76   */
    uint16_t unique_ips_24h;     /* FeatureVector.UniqueIPs24h                  engine.go:77   */
    uint8_t  tx_type;            /* IGS_TX_*                                                   */
    uint8_t  bits;               /* IGS_BIT_*                                                  */
    uint8_t  pad[2];
} igs_rule_row;

/* Result of scoring one transaction.  8 bytes.  Replaces the scalar part of ScoreResponse
 * (scoring/engine.go:56-64); rule_score is recomputed on the host from `flags`
 * (igs_rule_score_from_flags), reason strings from `flags` in emission order.                  */
typedef struct igs_result {
    float    ml_score; /* ScoreResponse.MLScore narrowed to the wire type (risk.proto:67)       */
    uint16_t flags;    /* IGS_R_* bits                                                          */
    uint8_t  score;    /* ScoreResponse.Score 0..100                                            */
    uint8_t  action;   /* IGS_ACTION_*                                                          */
} igs_result;

/* Compact transaction record for the fused path (assembly on device).  32 bytes.
 * Replaces the fields of ScoreRequest the reference actually reads (scoring/engine.go:40-53)
 * plus the host-side verdicts (IP intelligence :397, blacklist :470, HLL uniques
 * features/redis_store.go:80-81).                                                               */
typedef struct igs_tx {
    int64_t  amount;             /* cents                                                       */
    uint32_t account;            /* dense player index of this engine's shard                   */
    uint32_t ts;                 /* unix seconds; plays the role of time.Now() (engine.go:350)   */
    uint32_t seq_limit;          /* the query sees events of its account ordered before
                                    (ts, seq_limit); UINT32_MAX = every event with ts' <= ts     */
    uint16_t unique_devices_24h; /* host-supplied PFCOUNTs (redis_store.go:80-81); ignored when the */
    uint16_t unique_ips_24h;     /* history carries identifiers (igs_load_events_ids)            */
    uint8_t  tx_type;            /* IGS_TX_*                                                    */
    uint8_t  bits;               /* IGS_BIT_VPN|PROXY|TOR|BLACKLISTED|VALID                      */
    uint8_t  pad[6];
} igs_tx;

/* One player's hourly batch features: BatchFeatures, scoring/engine.go:127-140.  96 bytes.      */
typedef struct igs_player_row {
    int64_t  total_deposits;
    int64_t  total_withdrawals;
    int64_t  total_bets;
    int64_t  total_wins;
    int64_t  account_created_at;   /* unix seconds                                              */
    double   avg_bet_size;
    double   bonus_wager_complete;
    int32_t  deposit_count;
    int32_t  withdraw_count;
    int32_t  bet_count;
    int32_t  win_count;
    int32_t  bonus_claim_count;
    int32_t  reserved[5];
} igs_player_row;

/* Event history, struct of arrays (17 bytes per event).  Replaces the writes of
 * UpdateRealTimeFeatures, features/redis_store.go:119-168.  Arrival order = array order.        */
typedef struct igs_event_cols {
    const uint32_t* account;   /* dense player index                                            */
    const uint32_t* ts;        /* unix seconds                                                  */
    const int64_t*  amount;    /* cents                                                         */
    const uint8_t*  tx_type;   /* IGS_TX_*                                                      */
} igs_event_cols;

/* Velocity query: "what would GetRealTimeFeatures return for this account at time ts".          */
typedef struct igs_query {
    uint32_t account;
    uint32_t ts;
    uint32_t seq_limit;        /* see igs_tx.seq_limit                                          */
    uint32_t reserved;
} igs_query;

/* Per-type sliding window aggregate of the north-star extension (plain multiset semantics,
 * inclusive lower bound, prior events only).                                                   */
typedef struct igs_window {
    int64_t  sum;
    uint32_t count;
    uint32_t pad;
} igs_window;

/* Velocity block of one query.  Replaces RealTimeFeatures (features/redis_store.go:38-47 as
 * filled by GetRealTimeFeatures :60-116) and adds the extension windows.  128 bytes.           */
typedef struct igs_velocity {
    int64_t  tx_sum_1h;            /* Redis key tx_sum:1h — a chain sum, not a sliding window   */
    uint32_t tx_count_1m;          /* ZCOUNT [now-60,+inf)   distinct "ts:amount" members       */
    uint32_t tx_count_5m;          /* ZCOUNT [now-300,+inf)                                     */
    uint32_t tx_count_1h;          /* ZCOUNT [now-3600,+inf) after the 1h trim                  */
    uint32_t time_since_last_tx;   /* now - last_tx  iff key alive (7d), else 0                 */
    uint32_t session_duration;     /* now - session_start iff key alive (30m chain), else 0     */
    uint32_t n_prior;              /* events of the account ordered before the query            */
    igs_window win[2][3];          /* [IGS_TX_BET|IGS_TX_DEPOSIT][1m,1h,24h]                     */
} igs_velocity;

/* Identifiers — device id, browser fingerprint, IP (ScoreRequest.DeviceID / Fingerprint / IP,
 * scoring/engine.go:45-47; TransactionEvent.DeviceID / IP, features/redis_store.go:52-55) — cross
 * this ABI as 64-bit ids: igs_hash_id() of the string.  Id 0 is the empty string, which the
 * reference skips everywhere (`if event.DeviceID != ""` redis_store.go:141,148; :271-279).
 * Distinct strings are told apart by their 64-bit ids only (collision odds 2^-64 per pair).     */
typedef struct igs_event_ids {      /* columns parallel to igs_event_cols; a NULL column = all empty */
    const uint64_t* device;
    const uint64_t* ip;
} igs_event_ids;
typedef struct igs_tx_ids {         /* one per scored tx, parallel to igs_tx                      */
    uint64_t device;
    uint64_t fingerprint;
    uint64_t ip;
} igs_tx_ids;
/* RealTimeFeatures.UniqueDevices24h / UniqueIPs24h (features/redis_store.go:44-45).              */
typedef struct igs_uniques {
    uint32_t devices_24h;
    uint32_t ips_24h;
} igs_uniques;
/* The three blacklist sets, features/redis_store.go:245-247.                                     */
#define IGS_BL_DEVICE      0
#define IGS_BL_FINGERPRINT 1
#define IGS_BL_IP          2

/* Tree ensemble: n_trees complete depth-8 trees in implicit heap order (node i has children
 * 2i+1 / 2i+2; leaf j of the bottom level follows internal node 127+j/2).  Branch rule =
 * ONNX BRANCH_LEQ: go left iff x[feat] <= thr; NaN goes right.  margin = base_score +
 * leaves summed in tree-index order in fp32.  Stands in for fraud_model.onnx
 * (services/risk/cmd/main.go:62; ml/onnx_model.go:63-68,238).                                  */
typedef struct igs_gbdt_desc {
    uint32_t       n_trees;
    uint32_t       n_features;     /* <= IGS_NUM_FEATURES                                        */
    uint32_t       post_transform; /* IGS_POST_*                                                 */
    float          base_score;
    const uint8_t* feat;           /* n_trees * 255                                              */
    const float*   thr;            /* n_trees * 255                                              */
    const float*   leaf;           /* n_trees * 256                                              */
} igs_gbdt_desc;

/* LTV head: 256 -> 1024 -> 1024 -> 1024 -> 1, ReLU between layers, bf16 operands, fp32
 * accumulate.  Weights row-major [out][in] as raw bf16 bit patterns; biases fp32.  Stands in for
 * ltv_model.onnx (services/risk/cmd/main.go:63; ml/onnx_model.go:396-432).                    */
#define IGS_MLP_IN     256
#define IGS_MLP_HIDDEN 1024
typedef struct igs_mlp_desc {
    const uint16_t* w1; const float* b1; /* [1024][256], [1024]  */
    const uint16_t* w2; const float* b2; /* [1024][1024], [1024] */
    const uint16_t* w3; const float* b3; /* [1024][1024], [1024] */
    const uint16_t* w4; const float* b4; /* [1][1024], [1]       */
} igs_mlp_desc;

/* Per-player sideband of the LTV epilogue: the PlayerFeatures fields that the post-model steps
 * of LTVPredictor.Predict read (prediction/ltv.go:124-139,228-382).  80 bytes.                 */
typedef struct igs_ltv_row {
    double  sessions_per_week;
    double  deposit_frequency;
    double  total_deposits;
    double  total_withdrawals;
    double  net_revenue;
    double  bonus_conversion_rate;
    int32_t days_since_registration;
    int32_t days_since_last_deposit;
    int32_t days_since_last_bet;
    int32_t bet_count;
    int32_t support_tickets;
    int32_t bonuses_claimed;
    int32_t games_played;
    uint8_t push_enabled, email_opt_in, has_vip_manager, pad;
} igs_ltv_row;

/* LTVPrediction (prediction/ltv.go:26-35) without the uuid / timestamp.  32 bytes.              */
#define IGS_SEG_VIP 1
#define IGS_SEG_HIGH 2
#define IGS_SEG_MEDIUM 3
#define IGS_SEG_LOW 4
#define IGS_SEG_CHURNING 5       /* numbered as risk.proto:122-129 */
typedef struct igs_ltv_result {
    double  predicted_ltv;       /* adjusted LTV, ltv.go:127                                     */
    double  churn_risk;
    float   confidence;
    int32_t days_survival;
    uint8_t segment;             /* IGS_SEG_*                                                    */
    uint8_t next_best_action;    /* index into igs_ltv_action_name()                             */
    uint8_t pad[6];
} igs_ltv_result;

/* ml.LTVFeatures (ml/onnx_model.go:375-393): the 17 float32 inputs of ml.LTVModel, declaration
 * order.  68 bytes.                                                                             */
typedef struct igs_ltv_features {
    float days_since_registration, days_since_last_deposit, days_since_last_bet, total_active_days;
    float sessions_per_week, total_deposits, total_withdrawals, net_revenue, avg_deposit_amount;
    float deposit_frequency, total_bets, bet_count, win_rate, avg_bet_size, games_played;
    float bonuses_claimed, bonus_conversion_rate;
} igs_ltv_features;

/* ------------------------------------------------------------------------------------------ */
/* engine life cycle                                                                          */
/* ------------------------------------------------------------------------------------------ */

/* Replaces NewScoringEngine (scoring/engine.go:231-259) + ml.NewONNXModel (ml/onnx_model.go:44-82):
 * binds the engine to CUDA device `device`, creates IGS_NUM_SLOTS stream slots.  Fails with
 * IGS_ERR_CUDA when there is no usable device.                                                  */
int  igs_engine_create(int device, igs_engine** out);
/* Replaces ONNXModel.Close (ml/onnx_model.go:348-356).                                          */
void igs_engine_destroy(igs_engine* e);
/* Text of the last error on this engine (or of the last failed igs_engine_create when e==NULL). */
const char* igs_last_error(const igs_engine* e);

/* Pinned host memory for hot-path host buffers (cgo cannot hand Go memory to an async copy).   */
void* igs_alloc_pinned(size_t bytes);
void  igs_free_pinned(void* p);

/* Replaces the model file load of ml.NewONNXModel (ml/onnx_model.go:62-81).                      */
int igs_load_gbdt(igs_engine* e, const igs_gbdt_desc* d);
/* Replaces ml.NewLTVModel (ml/onnx_model.go:396-402).                                            */
int igs_load_mlp(igs_engine* e, const igs_mlp_desc* d);
/* Replaces FeatureStore.GetBatchFeatures (scoring/engine.go:110; "ClickHouse, computed hourly"):
 * the whole BatchFeatures table of this shard becomes device resident.                          */
int igs_load_players(igs_engine* e, const igs_player_row* rows, uint64_t n);
/* Replaces the Redis state written by UpdateRealTimeFeatures (features/redis_store.go:119-168):
 * sorts the events by (account, ts, arrival) and builds the per-account scan columns.
 * n_accounts = size of the dense account index space.                                          */
int igs_load_events(igs_engine* e, const igs_event_cols* ev, uint64_t n, uint32_t n_accounts);

/* igs_load_events / igs_load_events_ids with the columns already in device memory (a history restored
 * from a device-side snapshot, or produced by an upstream GPU stage): the structs hold device pointers
 * (d_ids may be NULL = no identifier columns).  The columns are copied; the caller keeps its buffers.
 * Synchronous like the host-buffer loads.  No counterpart in the reference.                         */
int igs_load_events_device(igs_engine* e, const igs_event_cols* d_ev, const igs_event_ids* d_ids, uint64_t n,
                           uint32_t n_accounts);

/* Replaces FeatureStore.UpdateRealTimeFeatures (scoring/engine.go:111; features/redis_store.go:119-168)
 * for a batch of completed transactions — what ScoringEngine.UpdateFeatures (scoring/engine.go:486-488)
 * does per "transaction.completed" event (services/risk/cmd/main.go:219-224): appends n events, in
 * arrival order after everything loaded or appended so far.  Every query afterwards answers exactly as
 * after igs_load_events of the concatenated history.
 * How: the raw columns stay on the device and only the new events cross PCIe.  The index built by the last
 * load stays as it is (the BASE); the events appended since form a second, small index (the DELTA) that is
 * rebuilt on every append — its chain sums, session starts and duplicate flags continue the base's — and the
 * lookup consults both.  When the delta outgrows max(threshold, base / 8) events the two are rebuilt into
 * one base (igs_set_compaction_threshold; default threshold 65 536).  Cost of an append = the delta's size,
 * not the history's; a compaction costs one full build.  Histories that carry identifiers
 * (igs_load_events_ids) still rebuild the whole index on every append.
 * Concurrency: an append builds into arrays no kernel is reading and then swaps them in, so scoring and
 * query calls on other slots / streams of the same engine may run WHILE an append is in progress: they see
 * the history either before or after the batch, never a mixture.  Appends themselves (and the loads,
 * igs_evict_events) are serialised by the engine.
 * Needs a prior igs_load_events (n = 0 is fine) that fixed n_accounts.  A batch with an out-of-range
 * account is rejected as a whole and the history is left as it was.                                  */
int igs_append_events(igs_engine* e, const igs_event_cols* ev, uint64_t n);
/* Allocates now what igs_append_events would otherwise allocate when it first needs it — room in the resident
 * columns for the events up to the next compaction, both sets of delta arrays, the spare base arrays a
 * compaction builds into — for batches of up to max_batch_events.  Optional; a live service calls it once
 * after igs_load_events so that no append pays for (and no scoring call stalls behind) a device allocation.
 * Costs about another index worth of device memory.  No counterpart in the reference.                  */
int igs_prepare_ingest(igs_engine* e, uint64_t max_batch_events);
/* The delta index is folded into the base when it holds more than max(min_delta_events, base / 8) events.
 * No counterpart in the reference (Redis updates in place).                                          */
int igs_set_compaction_threshold(igs_engine* e, uint64_t min_delta_events);

/* LATE EVENTS.  Redis applies every event when it arrives, with the event's own timestamp as "now"
 * (features/redis_store.go:121: ZADD score, :132 trim bound, :155 SET last_tx), so an event that arrives
 * after a newer event of the same account overwrites last_tx with an OLDER time, trims less, and renews
 * TTLs from the past.  The index has no arrival-order state: it files an event where its timestamp says,
 * i.e. it answers as if the events had been applied in (timestamp, arrival) order.  The two agree exactly
 * when, per account, timestamps never decrease in arrival order; they differ for a late event — one whose
 * timestamp is smaller than that of an event of the same account that arrived before it.
 *   igs_load_events* (a replayed log) accepts late events and orders them by timestamp.
 *   igs_append_events* (live ingest) by default REJECTS a batch that holds a late event (relative to the
 *   resident history or to earlier events of the batch) with IGS_ERR_INVALID and leaves the history as it
 *   was, so that the caller decides; with IGS_LATE_REORDER the batch is accepted and ordered by timestamp.
 * igs_history_info reports how many late pairs the resident history holds.                            */
#define IGS_LATE_REJECT  0
#define IGS_LATE_REORDER 1
int igs_set_late_policy(igs_engine* e, int policy);
typedef struct igs_history_info {
    uint64_t n_events;       /* resident events                                                       */
    uint64_t late_events;    /* neighbouring events of one account whose arrival order is the reverse
                                of their time order                                                   */
    uint32_t n_accounts;
    uint32_t has_identifiers;
    uint64_t delta_events;   /* of n_events: appended since the base index was last built (see igs_append_events) */
    uint64_t reserved[3];
} igs_history_info;
int igs_get_history_info(igs_engine* e, igs_history_info* out);

/* Replaces the key expiry Redis does on its own (the TTLs of features/redis_store.go:133,138,144,151,
 * 155,160): drops from the resident history every event that cannot influence any query at
 * t >= now, and refreshes the index.  What is dropped, per account: everything before the trailing
 * chain of events with consecutive gaps <= 86400 s (the longest-lived chain key; the windows reach
 * back 24 h too), and the whole account once now - (its last event) > 7 d (last_tx, the longest TTL).
 * Queries at t >= now return exactly what they returned before; queries in the past may not.  Events
 * dated after `now` (future-dated input, a caller whose clock is behind the ingest clock) are kept, and
 * so is everything a query between `now` and their timestamp can still see.  An
 * account that never pauses for a day keeps its whole history — as its tx_sum / unique-identifier keys
 * never expire in Redis either.  The survivors keep their arrival order but are renumbered: pass
 * seq_limit = UINT32_MAX in queries afterwards.  *n_evicted (optional) receives the number of events
 * dropped.  Like the loads, not concurrent with scoring calls on the same engine.                  */
int igs_evict_events(igs_engine* e, uint32_t now, uint64_t* n_evicted);

/* ------------------------------------------------------------------------------------------ */
/* scoring — replaces ScoringEngine.Score (scoring/engine.go:262-323) applied to a batch, i.e.   */
/* rpc ScoreBatch (proto/risk/v1/risk.proto:13,76-82) and ONNXModel.PredictBatch                 */
/* (ml/onnx_model.go:311-326)                                                                   */
/* ------------------------------------------------------------------------------------------ */

/* Path A: pre-assembled model input.  features = n x 64 fp32 rows, already normalised
 * (ml/onnx_model.go:169-184); rows = the rule sideband.  Host buffers; synchronous.
 * leaf_idx (optional, may be NULL) receives n x n_trees bytes: the leaf reached in every tree.  */
int igs_score_features(igs_engine* e, const igs_config* cfg, const float* features,
                       const igs_rule_row* rows, uint64_t n, igs_result* out, uint8_t* leaf_idx);
/* Same, asynchronous on slot `slot` (0..IGS_NUM_SLOTS-1); host buffers must be pinned and stay
 * valid until igs_wait(e, slot) returns.                                                        */
int igs_score_features_async(igs_engine* e, int slot, const igs_config* cfg, const float* features,
                             const igs_rule_row* rows, uint64_t n, igs_result* out);
int igs_wait(igs_engine* e, int slot);
/* Same on device-resident buffers; enqueues on `stream` (a cudaStream_t, NULL = default) and
 * returns without synchronising.                                                                */
int igs_score_features_device(igs_engine* e, const igs_config* cfg, const float* d_features,
                              const igs_rule_row* d_rows, uint64_t n, igs_result* d_out,
                              uint8_t* d_leaf_idx, void* stream);

/* Velocity only — replaces FeatureStore.GetRealTimeFeatures (features/redis_store.go:60-116)
 * for n queries against the loaded event history.                                              */
int igs_get_velocity(igs_engine* e, const igs_query* q, uint64_t n, igs_velocity* out);
int igs_get_velocity_device(igs_engine* e, const igs_query* d_q, uint64_t n, igs_velocity* d_out,
                        void* stream);

/* Path B, the fused path: compact tx records in, results out; velocity lookup, extractFeatures
 * (scoring/engine.go:326-417), the scoring->ml mapping + Normalize (ml/onnx_model.go:133-184),
 * applyRules (:420-483), the model and the blend (:284-310) all run on the device against the
 * loaded players / events / model.  features_out (optional, may be NULL) receives the n x 64
 * assembled model input, velocity_out (optional) the n velocity blocks.                         */
/* Host buffers.  Pageable memory is staged through the slot's device buffers in chunks (copy, kernel and copy
 * back overlap).  Large batches (>= 4 x 512 tx per SM) whose tx / out (/ ids) buffers come from igs_alloc_pinned
 * and that ask for the results only run as ONE kernel launch that starts at once: the records follow it into
 * device memory in 32 copies on a worker stream of the slot, the kernel waits for the chunk it is about to read,
 * and the results are written straight into the pinned result buffer.  (tx and out must stay untouched until the
 * call — or igs_wait for the asynchronous form — returns, as for every host-buffer call.)             */
int igs_score_batch(igs_engine* e, const igs_config* cfg, const igs_tx* tx, uint64_t n,
                    igs_result* out, float* features_out, igs_velocity* velocity_out);
int igs_score_batch_async(igs_engine* e, int slot, const igs_config* cfg, const igs_tx* tx,
                          uint64_t n, igs_result* out);
int igs_score_batch_device(igs_engine* e, const igs_config* cfg, const igs_tx* d_tx, uint64_t n,
                           igs_result* d_out, float* d_features_out, igs_velocity* d_velocity_out,
                           void* stream);

/* ------------------------------------------------------------------------------------------ */
/* identifiers on the device: exact unique devices / IPs per account and the      */
/* blacklist sets, so that R3, R4 and R8 need no host-side lookups                               */
/* ------------------------------------------------------------------------------------------ */

/* igs_load_events / igs_append_events with the events' device and IP identifiers — the PFADD +
 * EXPIRE 24h pair of UpdateRealTimeFeatures (features/redis_store.go:141-152).  A history loaded
 * through igs_load_events_ids "carries identifiers": from then on every scoring call on this
 * engine takes UniqueDevices24h / UniqueIPs24h from the history and ignores
 * igs_tx.unique_devices_24h / unique_ips_24h; plain igs_append_events on such a history appends
 * events without identifiers.  igs_append_events_ids on a history without identifiers fails with
 * IGS_ERR_STATE.
 * Semantics (per account, per key): an event with a non-empty identifier adds it to the key and
 * renews the key's 24h TTL, so the events that carried one form chains with consecutive gaps
 * <= 86400 s; a query at time t sees the distinct identifiers of the trailing chain of its prior
 * events iff t - (ts of the last of them) <= 86400, else 0.  Counts are EXACT where Redis answers
 * PFCOUNT from a HyperLogLog (0.81 % standard error): a deliberate deviation — the approximate
 * registers cannot be reproduced without Redis itself.  Counts saturate at 65 535.               */
int igs_load_events_ids(igs_engine* e, const igs_event_cols* ev, const igs_event_ids* ids, uint64_t n,
                        uint32_t n_accounts);
int igs_append_events_ids(igs_engine* e, const igs_event_cols* ev, const igs_event_ids* ids, uint64_t n);
/* The PFCOUNT pair of GetRealTimeFeatures (features/redis_store.go:80-81,102-107) for n queries.
 * Needs a history that carries identifiers.                                                     */
int igs_get_uniques(igs_engine* e, const igs_query* q, uint64_t n, igs_uniques* out);
int igs_get_uniques_device(igs_engine* e, const igs_query* d_q, uint64_t n, igs_uniques* d_out,
                           void* stream);

/* Replaces the blacklist sets of the feature store.  igs_blacklist_load = DEL + SADD of all n ids
 * (n = 0 empties the set); igs_blacklist_add = AddToBlacklist (features/redis_store.go:251-264)
 * for n ids.  kind = IGS_BL_*; an unknown kind fails with IGS_ERR_INVALID like the reference's
 * "unknown blacklist type" (:261).  Id 0 is ignored.  Like the other loads, not concurrent with
 * scoring calls on the same engine.                                                              */
int igs_blacklist_load(igs_engine* e, int kind, const uint64_t* ids, uint64_t n);
int igs_blacklist_add(igs_engine* e, int kind, const uint64_t* ids, uint64_t n);
/* Replaces Blacklist.Check (scoring/engine.go:174-176) = CheckBlacklist
 * (features/redis_store.go:267-293) for n identifier triples: out[i] = 1 iff any non-empty
 * identifier of ids[i] is in its set.                                                           */
int igs_blacklist_check(igs_engine* e, const igs_tx_ids* ids, uint64_t n, uint8_t* out);
int igs_blacklist_check_device(igs_engine* e, const igs_tx_ids* d_ids, uint64_t n, uint8_t* d_out,
                               void* stream);

/* igs_score_batch with rule 8 evaluated on the device (scoring/engine.go:468-475): the tx is
 * KNOWN_FRAUDSTER iff its IGS_BIT_BLACKLISTED is set or igs_blacklist_check(ids[i]) hits.  ids is
 * parallel to tx.  Everything else is igs_score_batch.                                          */
int igs_score_batch_ids(igs_engine* e, const igs_config* cfg, const igs_tx* tx, const igs_tx_ids* ids,
                        uint64_t n, igs_result* out, float* features_out, igs_velocity* velocity_out);
int igs_score_batch_ids_device(igs_engine* e, const igs_config* cfg, const igs_tx* d_tx,
                               const igs_tx_ids* d_ids, uint64_t n, igs_result* d_out,
                               float* d_features_out, igs_velocity* d_velocity_out, void* stream);

/* ------------------------------------------------------------------------------------------ */
/* linked accounts / bonus abuse — the one step of the path that crosses shards    */
/* ------------------------------------------------------------------------------------------ */
/* Serves RiskChecker.CheckBonusAbuse (services/bonus/internal/service/bonus_engine.go:139-141, called at
 * :268-275) and rpc CheckBonusAbuse (proto/risk/v1/risk.proto:20,135-145: is_abuser, abuse_score,
 * signals, linked_accounts).  The reference declares these interfaces and the MULTI_ACCOUNT reason
 * ("Linked to other accounts", risk.proto:274) but contains no implementation, so the semantics below
 * are DEFINED HERE (parity unpinned; the oracle restates them on the CPU):
 *   - two accounts are linked iff some device id occurs in the histories of both; a device seen on
 *     more than IGS_LINK_MAX_ACCOUNTS accounts is treated as public (a kiosk, an emulator image) and
 *     links nobody;
 *   - per account: n_linked = number of distinct linked accounts, linked[] = the three smallest of
 *     their global ids;
 *   - signals: IGS_ABUSE_BONUS_ONLY = FeatureVector.BonusOnlyPlayer (scoring/engine.go:384-386),
 *     IGS_ABUSE_MULTI_ACCOUNT = n_linked >= 1, IGS_ABUSE_ACCOUNT_RING = n_linked >= 3;
 *     abuse_score = (5*bonus_only + 3*multi_account + 2*account_ring) / 10; is_abuser = score >= 0.5.
 * Accounts live on the shard igs_shard_of picks, devices do not: the links are found in three stages
 * with an exchange between them.  A global account id is (shard << 32) | dense index.
 *   stage 1, every shard:  igs_links_pairs   -> (device, account) pairs of the local history, grouped
 *                                              by the shard that owns the device
 *                                              (owner = igs_shard_of(device, 0, n_shards))
 *   exchange 1:            all-to-all of the groups (the caller's collective: NCCL all_to_all over
 *                          NVLink in igaming-platform_b200/links.py; nothing to do for one shard)
 *   stage 2, device owner: igs_links_group   -> directed links (account, other), grouped by the shard
 *                                              that owns `account`
 *   exchange 2:            all-to-all of the groups
 *   stage 3, account owner: igs_links_finish -> the per-account table, resident on the device
 * Stage outputs are device buffers owned by the engine, valid until the next igs_links_* call;
 * counts[n_shards] (host) receives the group sizes, groups are contiguous in shard order.          */
#define IGS_LINK_MAX_ACCOUNTS 64
typedef struct igs_link_pair { uint64_t device; uint64_t account; } igs_link_pair;
typedef struct igs_link { uint64_t account; uint64_t other; } igs_link;
typedef struct igs_links_row {
    uint32_t n_linked;
    uint32_t reserved;
    uint64_t linked[3];        /* smallest global ids of the linked accounts, 0-filled              */
} igs_links_row;
#define IGS_ABUSE_BONUS_ONLY    (1u << 0)
#define IGS_ABUSE_MULTI_ACCOUNT (1u << 1)
#define IGS_ABUSE_ACCOUNT_RING  (1u << 2)
typedef struct igs_abuse {
    float    abuse_score;      /* CheckBonusAbuseResponse.abuse_score, 0..1                         */
    uint32_t n_linked;
    uint8_t  is_abuser;        /* RiskChecker.CheckBonusAbuse's bool                                */
    uint8_t  signals;          /* IGS_ABUSE_*                                                       */
    uint8_t  pad[6];
} igs_abuse;
int igs_links_pairs(igs_engine* e, uint32_t shard, uint32_t n_shards, const igs_link_pair** d_pairs,
                    uint64_t* counts);
int igs_links_group(igs_engine* e, const igs_link_pair* d_pairs, uint64_t n, uint32_t n_shards,
                    const igs_link** d_links, uint64_t* counts);
int igs_links_finish(igs_engine* e, const igs_link* d_links, uint64_t n, uint32_t shard);

/* The same job with the two exchanges fused into the producing kernels over peer memory (NVLink /
 * NVSwitch): every shard owns a receive WINDOW — one device allocation holding a cursor block, room
 * for cap_pairs pairs and cap_links links — that its peers map and write into directly.  Stage 1's
 * scatter and stage 2's emit claim slots in the destination's window with one remote atomic per block
 * and destination and store the records straight into it; nothing is staged, counted on the host or copied by
 * a collective.  No kernel waits for another GPU: the caller separates the steps with a barrier of
 * its own (torch.distributed.barrier in links.py), in this order on every shard:
 *     igs_links_window_reset -> barrier -> igs_links_pairs_peer -> barrier ->
 *     igs_links_group_peer   -> barrier -> igs_links_finish_peer
 * Every call returns after its kernels have completed.  A window that is too small loses nothing
 * silently: the cursor keeps counting, the consuming call fails with IGS_ERR_NOMEM and names the
 * capacity that would have sufficed.  No counterpart in the reference.
 * igs_links_window_create: allocates this shard's window and returns the 64-byte inter-process handle
 * (cudaIpcMemHandle_t) peers open; igs_links_window_open maps a peer's window from its handle (other
 * process / other GPU); igs_links_window_attach registers a window by address (this shard's own, or a
 * peer engine in the same process).  All n_shards windows, the own one included, must be attached /
 * opened before the first *_peer call; capacities must be the same on every shard.                  */
#define IGS_IPC_HANDLE_BYTES 64
int igs_links_window_create(igs_engine* e, uint64_t cap_pairs, uint64_t cap_links,
                            uint8_t handle_out[IGS_IPC_HANDLE_BYTES], void** d_window);
int igs_links_window_open(igs_engine* e, uint32_t peer_shard, const uint8_t handle[IGS_IPC_HANDLE_BYTES]);
int igs_links_window_attach(igs_engine* e, uint32_t peer_shard, void* d_window);
int igs_links_window_reset(igs_engine* e);
int igs_links_pairs_peer(igs_engine* e, uint32_t shard, uint32_t n_shards);
int igs_links_group_peer(igs_engine* e, uint32_t shard, uint32_t n_shards);
int igs_links_finish_peer(igs_engine* e, uint32_t shard);

/* Rows of the table built by igs_links_finish for n dense account indices of this shard (host
 * buffers).  An index beyond the table yields a zero row.                                          */
int igs_get_links(igs_engine* e, const uint32_t* accounts, uint64_t n, igs_links_row* out);
/* CheckBonusAbuse for n dense account indices (host buffers): needs igs_load_players and a links
 * table (igs_links_finish; without one every account has n_linked = 0).                            */
int igs_check_bonus_abuse(igs_engine* e, const uint32_t* accounts, uint64_t n, igs_abuse* out);
/* Signal strings in bit order: writes up to 3 pointers to static strings, returns how many.        */
int igs_abuse_signals(uint32_t signals, const char* out[3]);

/* Shard partition (device buffers): groups n tx records by shard = splitmix64(account) % n_shards
 * (n_shards <= 256), keeping arrival order inside a shard.  The reference has no batch partition;
 * the key is the one its feature store is addressed by (AccountID, features/redis_store.go:61,
 * scoring/engine.go:41) and the per-shard queues it would feed are QueueRiskScoring
 * (pkg/events/publisher.go:40).  d_tx_out[n] receives the grouped records, d_perm[n] (optional) the
 * original index of each output record, d_offsets[n_shards + 1] the first output position of each
 * shard.  Scratch comes from the engine (allocated on first use / growth).  Enqueues on `stream`.  */
int igs_partition_tx_device(igs_engine* e, const igs_tx* d_tx, uint64_t n, uint32_t n_shards,
                            igs_tx* d_tx_out, uint32_t* d_perm, uint32_t* d_offsets, void* stream);

/* The same partition for callers that must not allocate.  No counterpart in the reference.
 * Who they are: a shard's ingest thread that partitions batches while scoring calls of other slots
 * are in flight on the same GPU (igs_partition_tx_device obtains its counters with cudaMalloc on
 * first use and again on growth, and an allocation stalls everything running on the device), and
 * deployments that fix their GPU memory budget at start-up and treat a later allocation as a fault.
 * This variant calls no allocator at all: its counters are a 1 MiB array inside the library image
 * (IGS_PARTITION_STATIC_WORDS 32-bit words), resident on every GPU that loads the library, from
 * load until the process exits.
 *   - Limit: n_shards * ceil(n / 2048) <= IGS_PARTITION_STATIC_WORDS (8 shards: n <= 67 108 864;
 *     256 shards: n <= 2 097 152).  Larger requests fail with IGS_ERR_INVALID and a message that
 *     says so; nothing is launched.
 *   - The first call on a device looks up the array's device address (one runtime query, then
 *     cached).  Make that first call before starting any stream capture.
 *   - All calls on one device share the array, and it keeps whatever the previous call left.
 *     That is harmless for sequential calls (every counter a call reads, it has written first),
 *     but two calls whose kernels overlap on the GPU — different streams, different engines or
 *     threads on the same device — race on it and return garbage: order them yourself.
 * Arguments, outputs and stream semantics are those of igs_partition_tx_device.                   */
#define IGS_PARTITION_STATIC_WORDS (256u * 1024u)
int igs_partition_tx_static(igs_engine* e, const igs_tx* d_tx, uint64_t n, uint32_t n_shards,
                            igs_tx* d_tx_out, uint32_t* d_perm, uint32_t* d_offsets, void* stream);

/* LTV — replaces LTVPredictor.BatchPredict (prediction/ltv.go:385-398): x = n x 256 bf16 model
 * input (raw bit patterns), rows = epilogue sideband.  The MLP output replaces calculateLTV
 * (:155-178); churn / adjust / segment / survival / action / confidence follow :124-139.
 * ltv_raw (optional) receives the fp32 model output.                                           */
int igs_ltv_batch(igs_engine* e, const uint16_t* x, const igs_ltv_row* rows, uint64_t n,
                  igs_ltv_result* out, float* ltv_raw);
int igs_ltv_batch_device(igs_engine* e, const uint16_t* d_x, const igs_ltv_row* d_rows, uint64_t n,
                         igs_ltv_result* d_out, float* d_ltv_raw, void* stream);

/* Verify mode of the LTV model (the counterpart of leaf_idx in igs_score_features): runs the first
 * `layer` (1..3) layers on n <= 262144 rows and returns the n x 1024 bf16 activations, so a test can
 * check one GEMM at a time.  No reference counterpart (the reference's LTVModel holds session: nil,
 * ml/onnx_model.go:399).                                                                          */
int igs_ltv_hidden_device(igs_engine* e, const uint16_t* d_x, uint64_t n, int layer, uint16_t* d_hidden,
                          void* stream);

/* The reference's own LTV arithmetic, without any loaded model — what the LTV slot runs today.
 *
 * igs_ltv_predict_batch replaces LTVPredictor.BatchPredict (prediction/ltv.go:385-398) exactly as the
 * reference executes it: calculateLTV (:155-178) where igs_ltv_batch puts the MLP, then the same
 * churn / adjust / segment / survival / action / confidence steps (:124-139).  fp64 throughout,
 * bit-identical to the CPU restatement.  ltv_raw (optional) receives calculateLTV's value.
 *
 * igs_ltv_mock_predict replaces ml.LTVModel.Predict (ml/onnx_model.go:405-432, with
 * calculateEngagement :434-467 and calculateChurnRisk :469-490) for n feature rows: the model the
 * ml package returns while `session: nil // Use mock for now` (:399) — the LTV counterpart of
 * IGS_MODEL_MOCK.  ltv_out[i] = its float64 result.  Neither call needs igs_load_mlp.            */
int igs_ltv_predict_batch(igs_engine* e, const igs_ltv_row* rows, uint64_t n, igs_ltv_result* out,
                          double* ltv_raw);
int igs_ltv_predict_batch_device(igs_engine* e, const igs_ltv_row* d_rows, uint64_t n,
                                 igs_ltv_result* d_out, double* d_ltv_raw, void* stream);
int igs_ltv_mock_predict(igs_engine* e, const igs_ltv_features* f, uint64_t n, double* ltv_out);
int igs_ltv_mock_predict_device(igs_engine* e, const igs_ltv_features* d_f, uint64_t n, double* d_ltv_out,
                                void* stream);

/* ------------------------------------------------------------------------------------------ */
/* host-side helpers (pure C, no GPU)                                                          */
/* ------------------------------------------------------------------------------------------ */

/* min(100, sum of the weights of the rule bits in flags) — ScoreResponse.RuleScore,
 * scoring/engine.go:421-482.  ML_HIGH_RISK does not contribute.                                 */
int32_t igs_rule_score_from_flags(const igs_config* cfg, uint32_t flags);
/* Reason strings in the order applyRules emits them (R1..R8, then ML): writes up to
 * IGS_NUM_REASONS pointers to static strings into out[], returns how many.                      */
int igs_reason_codes(uint32_t flags, const char* out[IGS_NUM_REASONS]);
const char* igs_action_name(int action);       /* "approve" | "review" | "block"                */
const char* igs_ltv_action_name(int idx);      /* prediction/ltv.go:300-343 strings             */
const char* igs_segment_name(int seg);         /* prediction/ltv.go:17-23                       */
/* shard = splitmix64(lo ^ hi) % n_shards for a 16-byte account uuid (ScoreRequest.AccountID);
 * igs_shard_of(account, 0, n) is the shard igs_partition_tx_device assigns to a dense index.     */
uint32_t igs_shard_of(uint64_t uuid_lo, uint64_t uuid_hi, uint32_t n_shards);
/* Id of an identifier string (device id, fingerprint, IP text): FNV-1a 64 over the bytes followed
 * by the splitmix64 finaliser; len 0 -> 0, a non-empty string never maps to 0.  No reference
 * counterpart (Redis keys the strings themselves).                                              */
uint64_t igs_hash_id(const char* s, size_t len);
/* Number of CUDA kernels this library has launched in this process since it was loaded (all engines, all
 * devices) — instrumentation for benchmarks and tests; no reference counterpart.                  */
unsigned long long igs_kernel_launches(void);
/* Library / build identification, e.g. "igs 0.1 sm_100a".                                       */
const char* igs_version(void);

#ifdef __cplusplus
}
#endif
#endif /* IGS_H_ */
