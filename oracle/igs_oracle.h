/*
 * igs_oracle.h — CPU restatement of the reference's fraud-scoring path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under igaming-platform_b200/ may include, link or call
 * this.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs use it, and there only as the checker / the timed CPU baseline.
 *
 * What it restates (paths relative to the reference root, formeo/igaming-platform):
 *   services/risk/internal/scoring/engine.go      Score / extractFeatures / applyRules / blend
 *   services/risk/internal/ml/onnx_model.go       ml.FeatureVector / Normalize / mockPredict / LTVModel
 *   services/risk/internal/features/redis_store.go GetRealTimeFeatures / UpdateRealTimeFeatures
 *   services/risk/internal/prediction/ltv.go      LTVPredictor
 *
 * PINNING.  The reference has no tests, fixtures or golden vectors (zero *_test.go files), and
 * cannot be run here (no Go toolchain, no ONNX Runtime, no Redis; its packages do not even
 * type-check against each other).  The restatement is therefore pinned against known-answer
 * vectors hand-derived from the cited source lines (SURVEY.md §8c sets A, B, C; committed in
 * tests/golden/kat_*.json).
 * PARITY UNPINNED for everything behind the ONNX Runtime boundary: the reference would evaluate
 * fraud_model.onnx / ltv_model.onnx inside github.com/yalue/onnxruntime_go v1.6.0 (go.mod:14,
 * not vendored) and ships neither model.  The tree-ensemble and MLP evaluation orders, the
 * sigmoid and the bf16 rounding points below are DEFINED here (ONNX TreeEnsemble BRANCH_LEQ
 * semantics, leaves summed in tree order in fp32), not copied from anything.
 * Also unpinned: Redis PFCOUNT (HyperLogLog, approximate).  Where uniques are computed here
 * (orc_redis_update_ids, orc_uniques_*) they are EXACT distinct counts over the PFADD / EXPIRE-24h
 * key life time — a deliberate deviation: HLL registers cannot be restated bit for bit without
 * Redis's hash (MurmurHash64A with its seed) and sparse/dense encodings, none of which is under
 * /root/reference.
 */
#ifndef IGS_ORACLE_H_
#define IGS_ORACLE_H_

#include <stdint.h>
#include "../include/igs.h"

#ifdef __cplusplus
extern "C" {
#endif

/* scoring.FeatureVector, scoring/engine.go:67-105.  Go `int` is 64-bit on amd64. */
typedef struct orc_features {
    int64_t tx_count_1m, tx_count_5m, tx_count_1h;
    int64_t tx_sum_1h;
    double  tx_avg_1h;
    int64_t unique_devices_24h, unique_ips_24h, ip_country_changes, device_age;
    int64_t account_age;
    int64_t total_deposits, total_withdrawals, net_deposit;
    int64_t deposit_count, withdraw_count;
    int64_t time_since_last_tx, session_duration;
    double  avg_bet_size, win_rate;
    int32_t is_vpn, is_proxy, is_tor, disposable_email;
    int64_t bonus_claim_count;
    double  bonus_wager_rate;
    int32_t bonus_only_player;
    int32_t pad_;
} orc_features;

/* scoring.RealTimeFeatures, scoring/engine.go:115-124 (= features.RealTimeFeatures
 * redis_store.go:38-47). */
typedef struct orc_realtime {
    int64_t tx_count_1m, tx_count_5m, tx_count_1h;
    int64_t tx_sum_1h;
    int64_t unique_devices_24h, unique_ips_24h;
    int64_t last_tx_timestamp, session_start;
} orc_realtime;

typedef struct orc_score {
    int32_t  score;       /* ScoreResponse.Score  */
    int32_t  rule_score;  /* ScoreResponse.RuleScore */
    double   ml_score;    /* ScoreResponse.MLScore (f64 inside the engine) */
    uint32_t flags;       /* IGS_R_* */
    int32_t  action;      /* IGS_ACTION_* */
} orc_score;

/* ------------------------------------------------------------------------------------------ */
/* scoring/engine.go                                                                           */
/* ------------------------------------------------------------------------------------------ */
void orc_default_config(igs_config* cfg);                                /* :215-228, :246-257 */
/* extractFeatures :326-417.  rt / batch may be NULL (= that source failed -> zero features).   */
void orc_extract_features(const orc_realtime* rt, const igs_player_row* batch, int64_t now,
                          uint32_t ip_bits, orc_features* out);
/* applyRules :420-483.  Returns rule score; *flags receives the IGS_R_* rule bits.             */
int32_t orc_apply_rules(const igs_config* cfg, int64_t amount, int tx_type, const orc_features* f,
                        int blacklisted, uint32_t* flags);
/* blend + action :284-310.  has_model=0 is the `model == nil` arm (:276-288).                  */
void orc_blend(const igs_config* cfg, int32_t rule_score, uint32_t rule_flags, double ml,
               int has_model, orc_score* out);

/* ------------------------------------------------------------------------------------------ */
/* ml/onnx_model.go                                                                            */
/* ------------------------------------------------------------------------------------------ */
/* scoring.FeatureVector -> ml.FeatureVector (30 floats in ToSlice order :133-166).  The mapping
 * is not written in the reference (the two packages do not type-check); spec =
 * field-by-field name match, float32(int), bool -> 0|1, TxAmount = float32(req.Amount).        */
void   orc_to_ml(const orc_features* f, int64_t amount, int tx_type, float out30[30]);
void   orc_normalize(float v[30]);                                       /* :169-205           */
double orc_mock_predict(const float v[30]);                              /* :258-308           */
/* The deterministic fp32 sigmoid both sides implement (spec, see .c).                          */
float  orc_sigmoid(float margin);
/* Tree ensemble (spec in igs.h).  leaf_idx may be NULL; else n_trees bytes.                     */
float  orc_gbdt_margin(const igs_gbdt_desc* m, const float* x, uint8_t* leaf_idx);
float  orc_gbdt_predict(const igs_gbdt_desc* m, const float* x, uint8_t* leaf_idx);

/* ------------------------------------------------------------------------------------------ */
/* whole-row pipelines (what the GPU entry points are checked against)                          */
/* ------------------------------------------------------------------------------------------ */
/* Path A row: features already normalised (64 floats), rule sideband row.                      */
void orc_score_features_row(const igs_config* cfg, const igs_gbdt_desc* m, const float* x64,
                            const igs_rule_row* row, orc_score* out, uint8_t* leaf_idx);
/* Batch over n rows, `threads` host threads (static chunks).  out: igs_result per row.          */
void orc_score_features_batch(const igs_config* cfg, const igs_gbdt_desc* m, const float* x,
                              const igs_rule_row* rows, uint64_t n, igs_result* out,
                              uint8_t* leaf_idx, int threads);

/* ------------------------------------------------------------------------------------------ */
/* features/redis_store.go — literal emulation of the Redis commands on one account             */
/* ------------------------------------------------------------------------------------------ */
typedef struct orc_redis_account orc_redis_account;
orc_redis_account* orc_redis_new(void);
void orc_redis_free(orc_redis_account* a);
/* GetRealTimeFeatures :60-116 at wall clock `now` (unix s).  Uniques: exact set sizes.           */
void orc_redis_get(orc_redis_account* a, int64_t now, orc_realtime* out);
/* UpdateRealTimeFeatures :119-168 for an event at ts; wall clock == ts.                         */
void orc_redis_update(orc_redis_account* a, int64_t ts, int64_t amount);

/* Same with the event's device / IP identifiers (0 = empty string, skipped :141,:148).           */
void orc_redis_update_ids(orc_redis_account* a, int64_t ts, int64_t amount, uint64_t device,
                          uint64_t ip);
/* Closed form of one PFADD + EXPIRE-24h key against the account's n prior events.               */
uint32_t orc_uniques_closed_form(const uint32_t* ts, const uint64_t* id, uint32_t n_prior,
                                 uint32_t now);
void orc_uniques_batch(const igs_event_cols* ev, const igs_event_ids* ids, uint64_t n_events,
                       uint32_t n_accounts, const igs_query* q, uint64_t n_q, igs_uniques* out);
uint64_t orc_hash_id(const char* s, size_t len);                         /* spec of igs_hash_id */
/* Blacklist sets (SADD :251-264) and CheckBlacklist :267-293.                                   */
typedef struct orc_blacklist {
    const uint64_t* device; uint64_t n_device;
    const uint64_t* fingerprint; uint64_t n_fingerprint;
    const uint64_t* ip; uint64_t n_ip;
} orc_blacklist;
int orc_blacklist_check(const orc_blacklist* bl, const igs_tx_ids* id);

/* Closed form of the same + the extension windows, for one query against the account's events
 * already ordered by (ts, arrival): ts[], amount[], type[] of the n_prior events that precede
 * the query.  This is what the device kernel computes; tests prove it equal to the emulation.  */
void orc_velocity_closed_form(const uint32_t* ts, const int64_t* amount, const uint8_t* type,
                              uint32_t n_prior, uint32_t now, igs_velocity* out);
/* All queries against a whole history (any order; stable-sorted internally).                    */
void orc_velocity_batch(const igs_event_cols* ev, uint64_t n_events, uint32_t n_accounts,
                        const igs_query* q, uint64_t n_q, igs_velocity* out);

/* Path B row: compact tx + velocity block + player row -> 64-feature model input + score.      */
void orc_assemble(const igs_tx* tx, const igs_velocity* v, const igs_player_row* p,
                  orc_features* fv, float x64[64], igs_rule_row* row);
void orc_score_batch(const igs_config* cfg, const igs_gbdt_desc* m, const igs_player_row* players,
                     uint64_t n_players, const igs_event_cols* ev, uint64_t n_events,
                     const igs_tx* tx, uint64_t n, igs_result* out, float* features_out,
                     igs_velocity* velocity_out);

/* The same lookups from precomputed prefix columns and binary searches (O(log n) per query, like
 * Redis's ZCOUNT) — the timed CPU baseline of the fused workload; proven equal to the linear form.   */
typedef struct orc_fast_index orc_fast_index;
orc_fast_index* orc_fast_index_build(const igs_event_cols* ev, uint64_t n_events, uint32_t n_accounts);
void orc_fast_index_free(orc_fast_index* f);
void orc_fast_velocity(const orc_fast_index* f, uint32_t account, uint32_t ts, uint32_t seq_limit,
                       igs_velocity* out);
void orc_fast_velocity_batch(const orc_fast_index* f, const igs_query* q, uint64_t n_q, igs_velocity* out);
void orc_score_batch_indexed(const igs_config* cfg, const igs_gbdt_desc* m, const igs_player_row* players,
                             uint64_t n_players, const orc_fast_index* f, const igs_tx* tx, uint64_t n,
                             igs_result* out, int threads);

void orc_score_batch_ids(const igs_config* cfg, const igs_gbdt_desc* m, const igs_player_row* players,
                         uint64_t n_players, const igs_event_cols* ev, const igs_event_ids* ev_ids,
                         uint64_t n_events, const orc_blacklist* bl, const igs_tx* tx,
                         const igs_tx_ids* tx_ids, uint64_t n, igs_result* out, float* features_out);

/* Linked accounts over ALL shards at once (spec in igs.h, §8f-4): device[i] / account[i] = the
 * identifier and the global account id of event i; out[q] = row of query[q].  PARITY UNPINNED: the
 * reference declares CheckBonusAbuse (bonus_engine.go:139-141, risk.proto:135-145) without a body.  */
void orc_links(const uint64_t* device, const uint64_t* account, uint64_t n_events,
               const uint64_t* query, uint64_t n_q, igs_links_row* out);
void orc_check_bonus_abuse(const igs_player_row* p, const igs_links_row* l, igs_abuse* out);

/* ------------------------------------------------------------------------------------------ */
/* prediction/ltv.go + the MLP spec                                                            */
/* ------------------------------------------------------------------------------------------ */
double orc_ltv_engagement(const igs_ltv_row* f);                         /* :181-225           */
double orc_ltv_heuristic(const igs_ltv_row* f);                          /* calculateLTV :155-178 */
double orc_ltv_churn(const igs_ltv_row* f);                              /* :228-262           */
/* Everything after the model output, :124-139 (+ :265-382).                                    */
void   orc_ltv_epilogue(const igs_ltv_row* f, double ltv, igs_ltv_result* out);
/* ml.LTVModel (ml/onnx_model.go:405-490): Predict and its two helpers on ml.LTVFeatures.         */
double orc_ltv_model_engagement(const igs_ltv_features* f);              /* :434-467           */
double orc_ltv_model_churn(const igs_ltv_features* f);                   /* :469-490           */
double orc_ltv_model_predict(const igs_ltv_features* f);                 /* :405-432           */
float  orc_bf16_to_f32(uint16_t b);
uint16_t orc_f32_to_bf16(float f);                                       /* round-nearest-even */
/* MLP spec: bf16 operands, products exact, sums in fp64 then rounded once to fp32 per neuron,
 * +bias in fp32, ReLU, activations rounded to bf16 (layers 1-3); layer 4 output stays fp32.
 * hidden_out (optional): 3*1024 bf16 activations.                                              */
float  orc_mlp_forward(const igs_mlp_desc* m, const uint16_t* x256, uint16_t* hidden_out);
/* One layer only, from given bf16 inputs (teacher-forced check of a single GEMM).              */
void   orc_mlp_layer(const uint16_t* w, const float* b, int n_out, int n_in, const uint16_t* in,
                     float* pre_out);

#ifdef __cplusplus
}
#endif
#endif
