/*
 * igs_oracle.c — CPU restatement of the reference scoring path.  TEST INFRASTRUCTURE ONLY;
 * see igs_oracle.h for the scope, the pinning status ("parity unpinned" behind the ONNX Runtime
 * boundary) and who may call this.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math (see oracle/Makefile).  -ffp-contract=off
 * matters: Go on amd64 never fuses a*b+c, and engine.go:291-294 lands on different integers
 * if it is fused.
 */
#include "igs_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ========================================================================================== */
/* scoring/engine.go                                                                           */
/* ========================================================================================== */

/* DefaultConfig, scoring/engine.go:215-228; ruleWeights, :246-257. */
void orc_default_config(igs_config* c) {
    memset(c, 0, sizeof(*c));
    c->block_threshold = 80;
    c->review_threshold = 50;
    c->max_tx_per_minute = 10;
    c->max_tx_per_hour = 100;
    c->new_account_days = 7;
    c->large_deposit_amount = 100000;
    c->max_devices_per_day = 3;
    c->max_ips_per_day = 5;
    c->ml_weight = 0.6;
    c->rule_weight = 0.4;
    c->model_kind = IGS_MODEL_MOCK;
    const int32_t w[IGS_NUM_REASONS] = {20, 30, 25, 15, 20, 15, 50, 25, 20, 30};
    memcpy(c->reason_weight, w, sizeof(w));
}

/* extractFeatures, scoring/engine.go:326-417. */
void orc_extract_features(const orc_realtime* rt, const igs_player_row* b, int64_t now,
                          uint32_t ip_bits, orc_features* f) {
    memset(f, 0, sizeof(*f));
    if (rt) {                                         /* :343-354 */
        f->tx_count_1m = rt->tx_count_1m;
        f->tx_count_5m = rt->tx_count_5m;
        f->tx_count_1h = rt->tx_count_1h;
        f->tx_sum_1h = rt->tx_sum_1h;
        f->unique_devices_24h = rt->unique_devices_24h;
        f->unique_ips_24h = rt->unique_ips_24h;
        if (rt->last_tx_timestamp > 0) f->time_since_last_tx = now - rt->last_tx_timestamp;
        if (rt->session_start > 0) f->session_duration = now - rt->session_start;
    }
    if (b) {                                          /* :368-386 */
        f->total_deposits = b->total_deposits;
        f->total_withdrawals = b->total_withdrawals;
        f->net_deposit = b->total_deposits - b->total_withdrawals;
        f->deposit_count = b->deposit_count;
        f->withdraw_count = b->withdraw_count;
        f->avg_bet_size = b->avg_bet_size;
        /* int(time.Since(created).Hours() / 24): Duration.Hours() = float64(d/Hour) +
         * float64(d%Hour)/3.6e12, then /24, then truncation. */
        {
            int64_t d = now - b->account_created_at;            /* seconds */
            int64_t hour = d / 3600, rem = d % 3600;
            double hours = (double)hour + (double)(rem * 1000000000LL) / (60.0 * 60.0 * 1e9);
            f->account_age = (int64_t)(hours / 24.0);
        }
        f->bonus_claim_count = b->bonus_claim_count;
        f->bonus_wager_rate = b->bonus_wager_complete;
        if (b->bet_count > 0) f->win_rate = (double)b->win_count / (double)b->bet_count;
        if (b->bonus_claim_count > 3 && b->total_deposits < 5000) f->bonus_only_player = 1;
    }
    f->is_vpn = (ip_bits & IGS_BIT_VPN) != 0;        /* :403-405 */
    f->is_proxy = (ip_bits & IGS_BIT_PROXY) != 0;
    f->is_tor = (ip_bits & IGS_BIT_TOR) != 0;
    if (f->tx_count_1h > 0)                           /* :412-414 */
        f->tx_avg_1h = (double)f->tx_sum_1h / (double)f->tx_count_1h;
}

/* applyRules, scoring/engine.go:420-483.  All comparisons strict, rule order R1..R8. */
int32_t orc_apply_rules(const igs_config* c, int64_t amount, int tx_type, const orc_features* f,
                        int blacklisted, uint32_t* flags_out) {
    int32_t total = 0;
    uint32_t fl = 0;
    if (f->tx_count_1m > c->max_tx_per_minute) {                                     /* :425 */
        total += c->reason_weight[0]; fl |= IGS_R_HIGH_VELOCITY;
    }
    if (f->account_age < c->new_account_days && amount > c->large_deposit_amount) {   /* :431 */
        total += c->reason_weight[1]; fl |= IGS_R_NEW_ACCOUNT_LARGE_TX;
    }
    if (f->unique_devices_24h > c->max_devices_per_day) {                            /* :437 */
        total += c->reason_weight[3]; fl |= IGS_R_MULTIPLE_DEVICES;
    }
    if (f->unique_ips_24h > c->max_ips_per_day) {                                    /* :443 */
        total += c->reason_weight[2]; fl |= IGS_R_IP_COUNTRY_MISMATCH;
    }
    if (f->is_vpn || f->is_proxy || f->is_tor) {                                     /* :449 */
        total += c->reason_weight[5]; fl |= IGS_R_VPN_DETECTED;
    }
    if (f->time_since_last_tx < 300 && tx_type == IGS_TX_WITHDRAW) {                 /* :455 */
        /* int64: multiply, then truncating divide (:456) */
        if (f->deposit_count > 0 && f->total_withdrawals > f->total_deposits * 80 / 100) {
            total += c->reason_weight[7]; fl |= IGS_R_RAPID_DEPOSIT_WITHDRAW;
        }
    }
    if (f->bonus_only_player) {                                                      /* :463 */
        total += c->reason_weight[8]; fl |= IGS_R_BONUS_ABUSE;
    }
    if (blacklisted) {                                                               /* :469-475 */
        total += c->reason_weight[6]; fl |= IGS_R_KNOWN_FRAUDSTER;
    }
    if (total > 100) total = 100;                                                    /* :478 */
    *flags_out = fl;
    return total;
}

/* blend + action, scoring/engine.go:284-310. */
void orc_blend(const igs_config* c, int32_t rule_score, uint32_t rule_flags, double ml,
               int has_model, orc_score* out) {
    uint32_t fl = rule_flags;
    if (!has_model) ml = 0.0;                        /* var mlScore float64; model == nil :276 */
    else if (ml > 0.7) fl |= IGS_R_ML_HIGH_RISK;      /* :285-287 */
    /* :291-294, evaluated exactly as written, no contraction */
    volatile double a = c->rule_weight * (double)rule_score;
    volatile double m100 = ml * 100.0;
    volatile double b = c->ml_weight * m100;
    double s = a + b;
    int64_t fs = (int64_t)s;                         /* Go int(float64): truncation */
    if (fs > 100) fs = 100;                          /* :297-299 */
    int action;
    if (fs >= c->block_threshold) action = IGS_ACTION_BLOCK;          /* :303-310 */
    else if (fs >= c->review_threshold) action = IGS_ACTION_REVIEW;
    else action = IGS_ACTION_APPROVE;
    out->score = (int32_t)fs;
    out->rule_score = rule_score;
    out->ml_score = ml;
    out->flags = fl;
    out->action = action;
}

/* ========================================================================================== */
/* ml/onnx_model.go                                                                            */
/* ========================================================================================== */

/* ToSlice order, ml/onnx_model.go:133-166.  Mapping from scoring.FeatureVector = spec. */
void orc_to_ml(const orc_features* f, int64_t amount, int tx_type, float v[30]) {
    v[0] = (float)f->tx_count_1m;
    v[1] = (float)f->tx_count_5m;
    v[2] = (float)f->tx_count_1h;
    v[3] = (float)f->tx_sum_1h;
    v[4] = (float)f->tx_avg_1h;
    v[5] = (float)f->unique_devices_24h;
    v[6] = (float)f->unique_ips_24h;
    v[7] = (float)f->ip_country_changes;
    v[8] = (float)f->device_age;
    v[9] = (float)f->account_age;
    v[10] = (float)f->total_deposits;
    v[11] = (float)f->total_withdrawals;
    v[12] = (float)f->net_deposit;
    v[13] = (float)f->deposit_count;
    v[14] = (float)f->withdraw_count;
    v[15] = (float)f->time_since_last_tx;
    v[16] = (float)f->session_duration;
    v[17] = (float)f->avg_bet_size;
    v[18] = (float)f->win_rate;
    v[19] = f->is_vpn ? 1.0f : 0.0f;
    v[20] = f->is_proxy ? 1.0f : 0.0f;
    v[21] = f->is_tor ? 1.0f : 0.0f;
    v[22] = f->disposable_email ? 1.0f : 0.0f;
    v[23] = (float)f->bonus_claim_count;
    v[24] = (float)f->bonus_wager_rate;
    v[25] = f->bonus_only_player ? 1.0f : 0.0f;
    v[26] = (float)amount;
    v[27] = tx_type == IGS_TX_DEPOSIT ? 1.0f : 0.0f;
    v[28] = tx_type == IGS_TX_WITHDRAW ? 1.0f : 0.0f;
    v[29] = tx_type == IGS_TX_BET ? 1.0f : 0.0f;
}

/* logTransform, ml/onnx_model.go:186-195: log1p is the identity there (:193-195). */
static float log_transform(float x) {
    if (x <= 0) return 0;
    return (float)((double)x);
}
/* minMaxScale, :197-205 (min is always 0 at the call sites). */
static float min_max_scale(float x, float lo, float hi) {
    if (x < lo) return 0;
    if (x > hi) return 1;
    return (x - lo) / (hi - lo);
}
/* Normalize, :169-184. */
void orc_normalize(float v[30]) {
    v[3] = log_transform(v[3]);
    v[10] = log_transform(v[10]);
    v[11] = log_transform(v[11]);
    v[26] = log_transform(v[26]);
    v[0] = min_max_scale(v[0], 0, 20);
    v[1] = min_max_scale(v[1], 0, 50);
    v[2] = min_max_scale(v[2], 0, 200);
    v[5] = min_max_scale(v[5], 0, 10);
    v[6] = min_max_scale(v[6], 0, 20);
    v[9] = min_max_scale(v[9], 0, 365);
    v[15] = min_max_scale(v[15], 0, 86400);
}

/* mockPredict, :258-308.  float32 fields compared with constants converted to float32;
 * float64 accumulator, additions in source order. */
double orc_mock_predict(const float v[30]) {
    double score = 0;
    if (v[0] > 0.5f) score += 0.2;                                /* :262 */
    if (v[2] > 0.5f) score += 0.15;                               /* :265 */
    if (v[5] > 0.3f) score += 0.15;                               /* :270 */
    if (v[6] > 0.25f) score += 0.1;                               /* :273 */
    if (v[19] > 0 || v[20] > 0) score += 0.15;                    /* :278 */
    if (v[21] > 0) score += 0.25;                                 /* :281 */
    if (v[9] < 0.02f && v[26] > 0.5f) score += 0.2;               /* :286 */
    if (v[25] > 0) score += 0.15;                                 /* :291 */
    if (v[15] < 0.01f && v[28] > 0) {                             /* :296 */
        volatile float lim = v[10] * 0.8f;                        /* f32 multiply :297 */
        if (v[11] > lim) score += 0.2;
    }
    if (score > 1) score = 1;                                     /* :303 */
    return score;
}

/* Deterministic fp32 logistic (SPEC — parity unpinned; stands in for ONNX Runtime's).
 * Every step is one correctly-rounded IEEE-754 binary32 operation (mul, fma, add, div, rint),
 * so a CPU and a GPU evaluate it bit-identically:
 *   z = clamp(-m, -87, 88); n = rint(z*log2e); r = fma(n,-ln2hi,z); r = fma(n,-ln2lo,r);
 *   p = Horner degree 7 of e^r with fma; e = p * 2^n; return 1/(1+e).                         */
float orc_sigmoid(float m) {
    float z = -m;
    if (!(z >= -87.0f)) z = -87.0f;                   /* also catches NaN */
    if (z > 88.0f) z = 88.0f;
    volatile float t = z * 1.44269504088896341f;
    float n = rintf(t);
    float r = fmaf(n, -0.693145751953125f, z);
    r = fmaf(n, -1.42860682030941723e-06f, r);
    float p = 1.0f / 5040.0f;
    p = fmaf(p, r, 1.0f / 720.0f);
    p = fmaf(p, r, 1.0f / 120.0f);
    p = fmaf(p, r, 1.0f / 24.0f);
    p = fmaf(p, r, 1.0f / 6.0f);
    p = fmaf(p, r, 0.5f);
    p = fmaf(p, r, 1.0f);
    p = fmaf(p, r, 1.0f);
    union { uint32_t u; float f; } sc;
    sc.u = (uint32_t)((int32_t)n + 127) << 23;        /* 2^n, n in [-126,127] */
    volatile float e = p * sc.f;
    volatile float d = 1.0f + e;
    return 1.0f / d;
}

/* Tree ensemble (SPEC, igs.h igs_gbdt_desc). */
float orc_gbdt_margin(const igs_gbdt_desc* m, const float* x, uint8_t* leaf_idx) {
    float margin = m->base_score;   /* fp32 adds in tree order; -ffp-contract=off, SSE2 */
    for (uint32_t t = 0; t < m->n_trees; ++t) {
        const uint8_t* feat = m->feat + (size_t)t * IGS_TREE_NODES;
        const float* thr = m->thr + (size_t)t * IGS_TREE_NODES;
        uint32_t idx = 0;
        for (int d = 0; d < IGS_TREE_DEPTH; ++d) {
            float xv = x[feat[idx]];
            idx = 2 * idx + 1 + (xv <= thr[idx] ? 0u : 1u);   /* BRANCH_LEQ; NaN -> right */
        }
        uint32_t leaf = idx - IGS_TREE_NODES;
        if (leaf_idx) leaf_idx[t] = (uint8_t)leaf;
        margin = margin + m->leaf[(size_t)t * IGS_TREE_LEAVES + leaf];
    }
    return margin;
}

float orc_gbdt_predict(const igs_gbdt_desc* m, const float* x, uint8_t* leaf_idx) {
    float margin = orc_gbdt_margin(m, x, leaf_idx);
    float y = m->post_transform == IGS_POST_LOGISTIC ? orc_sigmoid(margin) : margin;
    if (y < 0) y = 0;                                 /* ml/onnx_model.go:247-252 */
    if (y > 1) y = 1;
    return y;
}

/* ========================================================================================== */
/* Path A: one pre-assembled row (per-request form; the CPU baseline loops this)                */
/* ========================================================================================== */
void orc_score_features_row(const igs_config* c, const igs_gbdt_desc* m, const float* x,
                            const igs_rule_row* r, orc_score* out, uint8_t* leaf_idx) {
    orc_features f;
    memset(&f, 0, sizeof(f));
    f.tx_count_1m = r->tx_count_1m;
    f.account_age = r->account_age_days;
    f.unique_devices_24h = r->unique_devices_24h;
    f.unique_ips_24h = r->unique_ips_24h;
    f.is_vpn = (r->bits & IGS_BIT_VPN) != 0;
    f.is_proxy = (r->bits & IGS_BIT_PROXY) != 0;
    f.is_tor = (r->bits & IGS_BIT_TOR) != 0;
    f.time_since_last_tx = r->time_since_last_tx;
    f.deposit_count = r->deposit_count;
    f.total_deposits = r->total_deposits;
    f.total_withdrawals = r->total_withdrawals;
    f.bonus_only_player = (r->bits & IGS_BIT_BONUS_ONLY) != 0;
    uint32_t flags;
    int32_t rule = orc_apply_rules(c, r->amount, r->tx_type, &f, (r->bits & IGS_BIT_BLACKLISTED) != 0,
                                   &flags);
    double ml = 0;
    int has_model = c->model_kind != IGS_MODEL_NONE;
    if (has_model) {
        if (!(r->bits & IGS_BIT_VALID)) {
            ml = 0.5;                                 /* scoring/engine.go:279-282 */
            if (leaf_idx && m && c->model_kind == IGS_MODEL_GBDT) memset(leaf_idx, 0, m->n_trees);
        } else if (c->model_kind == IGS_MODEL_MOCK) {
            ml = orc_mock_predict(x);
        } else {
            ml = (double)orc_gbdt_predict(m, x, leaf_idx);   /* float64(result) :254 */
        }
    }
    orc_blend(c, rule, flags, ml, has_model, out);
}

static void to_result(const orc_score* s, igs_result* r) {
    r->ml_score = (float)s->ml_score;
    r->flags = (uint16_t)s->flags;
    r->score = (uint8_t)s->score;
    r->action = (uint8_t)s->action;
}

typedef struct {
    const igs_config* c; const igs_gbdt_desc* m; const float* x; const igs_rule_row* rows;
    uint64_t lo, hi; igs_result* out; uint8_t* leaf;
} sf_job;

static void* sf_worker(void* p) {
    sf_job* j = (sf_job*)p;
    uint32_t nt = j->m ? j->m->n_trees : 0;
    for (uint64_t i = j->lo; i < j->hi; ++i) {
        orc_score s;
        orc_score_features_row(j->c, j->m, j->x + i * IGS_NUM_FEATURES, j->rows + i, &s,
                               j->leaf ? j->leaf + i * nt : NULL);
        to_result(&s, j->out + i);
    }
    return NULL;
}

void orc_score_features_batch(const igs_config* c, const igs_gbdt_desc* m, const float* x,
                              const igs_rule_row* rows, uint64_t n, igs_result* out,
                              uint8_t* leaf_idx, int threads) {
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t th[256];
    sf_job jobs[256];
    uint64_t chunk = (n + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
        uint64_t lo = (uint64_t)t * chunk, hi = lo + chunk;
        if (lo > n) lo = n;
        if (hi > n) hi = n;
        jobs[t] = (sf_job){c, m, x, rows, lo, hi, out, leaf_idx};
        if (threads == 1) sf_worker(&jobs[t]);
        else pthread_create(&th[t], NULL, sf_worker, &jobs[t]);
    }
    if (threads > 1)
        for (int t = 0; t < threads; ++t) pthread_join(th[t], NULL);
}

/* ========================================================================================== */
/* features/redis_store.go — literal emulation of the Redis commands for one account            */
/* ========================================================================================== */
struct orc_redis_account {
    /* ZSET features:<id>:tx_history — members "ts:amount", score ts (:127-130) */
    int64_t* z_ts; int64_t* z_amt; size_t z_n, z_cap;
    int64_t z_expire;            /* EXPIRE 2h :133; <0 = key absent */
    int64_t sum, sum_expire;     /* tx_sum:1h  INCRBY + EXPIRE 1h :136-138 */
    int64_t last, last_expire;   /* last_tx    SET EX 7d :155 */
    int64_t sess, sess_expire;   /* session_start SETNX + EXPIRE 30m :158-160 */
    /* unique_devices / unique_ips: PFADD + EXPIRE 24h :141-152.  Exact sets stand in for the
     * HyperLogLog registers (PFCOUNT is approximate; the deviation is documented in the header). */
    uint64_t* dev; size_t dev_n, dev_cap; int64_t dev_expire;
    uint64_t* ip;  size_t ip_n,  ip_cap;  int64_t ip_expire;
};

orc_redis_account* orc_redis_new(void) {
    orc_redis_account* a = (orc_redis_account*)calloc(1, sizeof(*a));
    a->z_expire = a->sum_expire = a->last_expire = a->sess_expire = -1;
    a->dev_expire = a->ip_expire = -1;
    return a;
}
void orc_redis_free(orc_redis_account* a) {
    if (!a) return;
    free(a->z_ts); free(a->z_amt); free(a->dev); free(a->ip); free(a);
}
/* Redis expires a key when now > when (keyIsExpired); at now == when it is still alive. */
static int alive(int64_t expire, int64_t now) { return expire >= 0 && !(now > expire); }

static void redis_expire_keys(orc_redis_account* a, int64_t now) {
    if (!alive(a->z_expire, now)) { a->z_n = 0; a->z_expire = -1; }
    if (!alive(a->sum_expire, now)) { a->sum = 0; a->sum_expire = -1; }
    if (!alive(a->last_expire, now)) { a->last = 0; a->last_expire = -1; }
    if (!alive(a->sess_expire, now)) { a->sess = 0; a->sess_expire = -1; }
    if (!alive(a->dev_expire, now)) { a->dev_n = 0; a->dev_expire = -1; }
    if (!alive(a->ip_expire, now)) { a->ip_n = 0; a->ip_expire = -1; }
}

/* GetRealTimeFeatures, features/redis_store.go:60-116. */
void orc_redis_get(orc_redis_account* a, int64_t now, orc_realtime* out) {
    memset(out, 0, sizeof(*out));
    redis_expire_keys(a, now);
    int64_t min1 = now - 60, min5 = now - 300, hour = now - 3600;     /* :67-69 */
    for (size_t i = 0; i < a->z_n; ++i) {                             /* ZCOUNT [x,+inf) :72-74 */
        if (a->z_ts[i] >= min1) out->tx_count_1m++;
        if (a->z_ts[i] >= min5) out->tx_count_5m++;
        if (a->z_ts[i] >= hour) out->tx_count_1h++;
    }
    if (a->sum_expire >= 0) out->tx_sum_1h = a->sum;                  /* :99-101 */
    if (a->last_expire >= 0) out->last_tx_timestamp = a->last;        /* :108-110 */
    if (a->sess_expire >= 0) out->session_start = a->sess;            /* :111-113 */
    out->unique_devices_24h = (int64_t)a->dev_n;                      /* PFCOUNT :80, :102-104 */
    out->unique_ips_24h = (int64_t)a->ip_n;                           /* PFCOUNT :81, :105-107 */
}

/* UpdateRealTimeFeatures, features/redis_store.go:119-168; wall clock == event ts. */
void orc_redis_update(orc_redis_account* a, int64_t ts, int64_t amount) {
    redis_expire_keys(a, ts);
    /* ZADD member "ts:amount" :127-130 — an existing member is not duplicated */
    int found = 0;
    for (size_t i = 0; i < a->z_n; ++i)
        if (a->z_ts[i] == ts && a->z_amt[i] == amount) { found = 1; break; }
    if (!found) {
        if (a->z_n == a->z_cap) {
            a->z_cap = a->z_cap ? a->z_cap * 2 : 16;
            a->z_ts = (int64_t*)realloc(a->z_ts, a->z_cap * sizeof(int64_t));
            a->z_amt = (int64_t*)realloc(a->z_amt, a->z_cap * sizeof(int64_t));
        }
        a->z_ts[a->z_n] = ts; a->z_amt[a->z_n] = amount; a->z_n++;
    }
    /* ZREMRANGEBYSCORE -inf .. ts-3600 (inclusive) :132 */
    size_t w = 0;
    for (size_t i = 0; i < a->z_n; ++i)
        if (!(a->z_ts[i] <= ts - 3600)) { a->z_ts[w] = a->z_ts[i]; a->z_amt[w] = a->z_amt[i]; w++; }
    a->z_n = w;
    a->z_expire = ts + 7200;                                          /* :133 */
    a->sum += amount; a->sum_expire = ts + 3600;                      /* :137-138 */
    a->last = ts; a->last_expire = ts + 7 * 24 * 3600;                /* :155 */
    if (a->sess_expire < 0) a->sess = ts;                             /* SETNX :159 */
    a->sess_expire = ts + 1800;                                       /* :160 */
}

static void set_add(uint64_t** v, size_t* n, size_t* cap, uint64_t id) {
    for (size_t i = 0; i < *n; ++i)
        if ((*v)[i] == id) return;
    if (*n == *cap) {
        *cap = *cap ? *cap * 2 : 8;
        *v = (uint64_t*)realloc(*v, *cap * sizeof(uint64_t));
    }
    (*v)[(*n)++] = id;
}

/* UpdateRealTimeFeatures with the event's identifiers: id 0 = empty string, which the reference
 * skips (`if event.DeviceID != ""` :141, `if event.IP != ""` :148). */
void orc_redis_update_ids(orc_redis_account* a, int64_t ts, int64_t amount, uint64_t device,
                          uint64_t ip) {
    orc_redis_update(a, ts, amount);
    if (device) { set_add(&a->dev, &a->dev_n, &a->dev_cap, device); a->dev_expire = ts + 86400; } /* :143-144 */
    if (ip) { set_add(&a->ip, &a->ip_n, &a->ip_cap, ip); a->ip_expire = ts + 86400; }             /* :150-151 */
}

static int u64_cmp(const void* a, const void* b) {
    const uint64_t x = *(const uint64_t*)a, y = *(const uint64_t*)b;
    return x < y ? -1 : x > y;
}

/* Closed form of one PFADD/EXPIRE-24h key: the events that carried an identifier form chains with
 * consecutive gaps <= 86400 (each PFADD refreshes the TTL); the key holds the distinct identifiers
 * of the trailing chain and is alive iff now - (ts of the last such event) <= 86400. */
uint32_t orc_uniques_closed_form(const uint32_t* ts, const uint64_t* id, uint32_t n, uint32_t now_u) {
    const int64_t now = now_u;
    uint32_t j = n;
    while (j > 0 && id[j - 1] == 0) --j;
    if (j == 0) return 0;
    int64_t last = ts[j - 1];
    if (now - last > 86400) return 0;
    uint64_t* chain = (uint64_t*)malloc((size_t)j * sizeof(uint64_t));
    size_t m = 0;
    for (uint32_t k = j; k-- > 0;) {
        if (id[k] == 0) continue;
        if (last - (int64_t)ts[k] > 86400) break;
        chain[m++] = id[k];
        last = ts[k];
    }
    qsort(chain, m, sizeof(uint64_t), u64_cmp);
    uint32_t distinct = 0;
    for (size_t k = 0; k < m; ++k)
        if (k == 0 || chain[k] != chain[k - 1]) ++distinct;
    free(chain);
    return distinct;
}

/* Closed form.  Prior events ordered by (ts, arrival). */
void orc_velocity_closed_form(const uint32_t* ts, const int64_t* amount, const uint8_t* type,
                              uint32_t n, uint32_t now_u, igs_velocity* v) {
    memset(v, 0, sizeof(*v));
    v->n_prior = n;
    if (n == 0) return;
    const int64_t now = now_u;
    const int64_t L = ts[n - 1];
    /* members alive in the ZSET: score > L-3600 (last trim), distinct (ts, amount) */
    for (uint32_t j = 0; j < n; ++j) {
        int64_t t = ts[j];
        if (t > L - 3600) {
            int dup = 0;
            for (uint32_t k = j; k-- > 0;) {
                if (ts[k] != ts[j]) break;
                if (amount[k] == amount[j]) { dup = 1; break; }
            }
            if (!dup) {
                if (t >= now - 60) v->tx_count_1m++;
                if (t >= now - 300) v->tx_count_5m++;
                if (t >= now - 3600) v->tx_count_1h++;
            }
        }
        if (type[j] == IGS_TX_BET || type[j] == IGS_TX_DEPOSIT) {
            igs_window* w = v->win[type[j]];
            if (t >= now - 60) { w[0].count++; w[0].sum += amount[j]; }
            if (t >= now - 3600) { w[1].count++; w[1].sum += amount[j]; }
            if (t >= now - 86400) { w[2].count++; w[2].sum += amount[j]; }
        }
    }
    if (now - L <= 3600) {                       /* tx_sum:1h chain, gaps <= 3600 */
        int64_t s = amount[n - 1];
        for (uint32_t j = n - 1; j > 0 && (int64_t)ts[j] - (int64_t)ts[j - 1] <= 3600; --j)
            s += amount[j - 1];
        v->tx_sum_1h = s;
    }
    if (now - L <= 7 * 24 * 3600) v->time_since_last_tx = (uint32_t)(now - L);
    if (now - L <= 1800) {                       /* session chain, gaps <= 1800 */
        uint32_t j = n - 1;
        while (j > 0 && (int64_t)ts[j] - (int64_t)ts[j - 1] <= 1800) --j;
        v->session_duration = (uint32_t)(now - ts[j]);
    }
}

typedef struct { uint32_t account, ts, seq; } ev_key;
static int ev_cmp(const void* a, const void* b) {
    const ev_key* x = (const ev_key*)a; const ev_key* y = (const ev_key*)b;
    if (x->account != y->account) return x->account < y->account ? -1 : 1;
    if (x->ts != y->ts) return x->ts < y->ts ? -1 : 1;
    if (x->seq != y->seq) return x->seq < y->seq ? -1 : 1;
    return 0;
}

typedef struct {
    ev_key* keys; uint64_t* seg;          /* seg[a]..seg[a+1] = events of account a */
    uint32_t* ts; int64_t* amt; uint8_t* ty; uint32_t* seq;
    uint64_t* dev; uint64_t* ip;          /* NULL when the history has no identifier columns */
    uint64_t n; uint32_t n_accounts;
} ev_index;

static void ev_index_build(ev_index* ix, const igs_event_cols* ev, uint64_t n, uint32_t n_accounts) {
    ix->n = n; ix->n_accounts = n_accounts;
    ix->dev = ix->ip = NULL;
    ix->keys = (ev_key*)malloc((n ? n : 1) * sizeof(ev_key));
    for (uint64_t i = 0; i < n; ++i) ix->keys[i] = (ev_key){ev->account[i], ev->ts[i], (uint32_t)i};
    qsort(ix->keys, n, sizeof(ev_key), ev_cmp);
    ix->ts = (uint32_t*)malloc((n ? n : 1) * 4);
    ix->amt = (int64_t*)malloc((n ? n : 1) * 8);
    ix->ty = (uint8_t*)malloc(n ? n : 1);
    ix->seq = (uint32_t*)malloc((n ? n : 1) * 4);
    ix->seg = (uint64_t*)calloc((size_t)n_accounts + 1, sizeof(uint64_t));
    for (uint64_t i = 0; i < n; ++i) {
        uint32_t s = ix->keys[i].seq;
        ix->ts[i] = ev->ts[s]; ix->amt[i] = ev->amount[s]; ix->ty[i] = ev->tx_type[s]; ix->seq[i] = s;
        if (ix->keys[i].account < n_accounts) ix->seg[ix->keys[i].account + 1]++;
    }
    for (uint32_t a = 0; a < n_accounts; ++a) ix->seg[a + 1] += ix->seg[a];
}
static void ev_index_add_ids(ev_index* ix, const igs_event_ids* ids) {
    const uint64_t n = ix->n;
    ix->dev = (uint64_t*)calloc(n ? n : 1, 8);
    ix->ip = (uint64_t*)calloc(n ? n : 1, 8);
    for (uint64_t i = 0; i < n; ++i) {
        if (ids->device) ix->dev[i] = ids->device[ix->seq[i]];
        if (ids->ip) ix->ip[i] = ids->ip[ix->seq[i]];
    }
}
static void ev_index_free(ev_index* ix) {
    free(ix->keys); free(ix->seg); free(ix->ts); free(ix->amt); free(ix->ty); free(ix->seq);
    free(ix->dev); free(ix->ip);
}
/* number of events of `account` ordered before (ts, seq_limit); *start = first sorted position */
static uint64_t ev_index_prior(const ev_index* ix, uint32_t account, uint32_t ts, uint32_t seq_limit,
                               uint64_t* start) {
    *start = 0;
    if (account >= ix->n_accounts) return 0;
    uint64_t s = ix->seg[account], e = ix->seg[account + 1];
    uint64_t u = s;
    while (u < e && (ix->ts[u] < ts || (ix->ts[u] == ts && ix->seq[u] < seq_limit))) ++u;
    *start = s;
    return u - s;
}
static void ev_index_uniques(const ev_index* ix, uint32_t account, uint32_t ts, uint32_t seq_limit,
                             igs_uniques* out) {
    uint64_t s;
    const uint32_t n = (uint32_t)ev_index_prior(ix, account, ts, seq_limit, &s);
    out->devices_24h = ix->dev ? orc_uniques_closed_form(ix->ts + s, ix->dev + s, n, ts) : 0;
    out->ips_24h = ix->ip ? orc_uniques_closed_form(ix->ts + s, ix->ip + s, n, ts) : 0;
}
static void ev_index_query(const ev_index* ix, uint32_t account, uint32_t ts, uint32_t seq_limit,
                           igs_velocity* out) {
    if (account >= ix->n_accounts) { memset(out, 0, sizeof(*out)); return; }
    uint64_t s = ix->seg[account], e = ix->seg[account + 1];
    uint64_t u = s;
    while (u < e && (ix->ts[u] < ts || (ix->ts[u] == ts && ix->seq[u] < seq_limit))) ++u;
    orc_velocity_closed_form(ix->ts + s, ix->amt + s, ix->ty + s, (uint32_t)(u - s), ts, out);
}

void orc_velocity_batch(const igs_event_cols* ev, uint64_t n_events, uint32_t n_accounts,
                        const igs_query* q, uint64_t n_q, igs_velocity* out) {
    ev_index ix;
    ev_index_build(&ix, ev, n_events, n_accounts);
    for (uint64_t i = 0; i < n_q; ++i)
        ev_index_query(&ix, q[i].account, q[i].ts, q[i].seq_limit, out + i);
    ev_index_free(&ix);
}

void orc_uniques_batch(const igs_event_cols* ev, const igs_event_ids* ids, uint64_t n_events,
                       uint32_t n_accounts, const igs_query* q, uint64_t n_q, igs_uniques* out) {
    ev_index ix;
    ev_index_build(&ix, ev, n_events, n_accounts);
    ev_index_add_ids(&ix, ids);
    for (uint64_t i = 0; i < n_q; ++i)
        ev_index_uniques(&ix, q[i].account, q[i].ts, q[i].seq_limit, out + i);
    ev_index_free(&ix);
}

/* The id of an identifier string (spec of igs_hash_id, include/igs.h): FNV-1a 64 over the bytes,
 * then the splitmix64 finaliser; "" -> 0 and a non-empty string never maps to 0. */
uint64_t orc_hash_id(const char* s, size_t len) {
    if (len == 0) return 0;
    uint64_t h = 0xcbf29ce484222325ull;
    for (size_t i = 0; i < len; ++i) { h ^= (uint8_t)s[i]; h *= 0x100000001b3ull; }
    h ^= h >> 30; h *= 0xbf58476d1ce4e5b9ull;
    h ^= h >> 27; h *= 0x94d049bb133111ebull;
    h ^= h >> 31;
    return h ? h : 1;
}

/* CheckBlacklist, features/redis_store.go:267-293: SISMEMBER per non-empty identifier, any hit. */
static int set_has(const uint64_t* v, uint64_t n, uint64_t id) {
    if (!id) return 0;                               /* `if deviceID != ""` :271, :274, :277 */
    for (uint64_t i = 0; i < n; ++i)
        if (v[i] == id) return 1;
    return 0;
}
int orc_blacklist_check(const orc_blacklist* bl, const igs_tx_ids* id) {
    return set_has(bl->device, bl->n_device, id->device) ||
           set_has(bl->fingerprint, bl->n_fingerprint, id->fingerprint) ||
           set_has(bl->ip, bl->n_ip, id->ip);
}

/* ========================================================================================== */
/* Path B: assemble (extractFeatures + scoring->ml mapping + Normalize) and score                */
/* ========================================================================================== */
void orc_assemble(const igs_tx* tx, const igs_velocity* v, const igs_player_row* p,
                  orc_features* fv, float x[64], igs_rule_row* row) {
    orc_realtime rt;
    memset(&rt, 0, sizeof(rt));
    const int64_t now = tx->ts;
    rt.tx_count_1m = v->tx_count_1m;
    rt.tx_count_5m = v->tx_count_5m;
    rt.tx_count_1h = v->tx_count_1h;
    rt.tx_sum_1h = v->tx_sum_1h;
    rt.unique_devices_24h = tx->unique_devices_24h;
    rt.unique_ips_24h = tx->unique_ips_24h;
    /* the velocity block already carries now-last / now-start (0 when the key is dead) */
    rt.last_tx_timestamp = v->time_since_last_tx ? now - v->time_since_last_tx : 0;
    rt.session_start = v->session_duration ? now - v->session_duration : 0;
    orc_extract_features(&rt, p, now, tx->bits, fv);
    float m[30];
    orc_to_ml(fv, tx->amount, tx->tx_type, m);
    orc_normalize(m);
    memset(x, 0, 64 * sizeof(float));
    memcpy(x, m, sizeof(m));
    int k = 30;                                   /* extension block, idx 30..41 */
    for (int ty = 0; ty < 2; ++ty)
        for (int w = 0; w < 3; ++w) {
            x[k++] = (float)v->win[ty][w].count;
            x[k++] = (float)v->win[ty][w].sum;
        }
    memset(row, 0, sizeof(*row));
    row->amount = tx->amount;
    row->total_deposits = fv->total_deposits;
    row->total_withdrawals = fv->total_withdrawals;
    row->tx_count_1m = (uint32_t)fv->tx_count_1m;
    row->account_age_days = (int32_t)fv->account_age;
    row->time_since_last_tx = (uint32_t)fv->time_since_last_tx;
    row->deposit_count = (uint32_t)fv->deposit_count;
    row->unique_devices_24h = tx->unique_devices_24h;
    row->unique_ips_24h = tx->unique_ips_24h;
    row->tx_type = tx->tx_type;
    row->bits = (uint8_t)((tx->bits & (IGS_BIT_VPN | IGS_BIT_PROXY | IGS_BIT_TOR | IGS_BIT_BLACKLISTED |
                                       IGS_BIT_VALID)) |
                          (fv->bonus_only_player ? IGS_BIT_BONUS_ONLY : 0));
}

void orc_score_batch(const igs_config* c, const igs_gbdt_desc* m, const igs_player_row* players,
                     uint64_t n_players, const igs_event_cols* ev, uint64_t n_events,
                     const igs_tx* tx, uint64_t n, igs_result* out, float* features_out,
                     igs_velocity* velocity_out) {
    ev_index ix;
    ev_index_build(&ix, ev, n_events, (uint32_t)n_players);
    for (uint64_t i = 0; i < n; ++i) {
        igs_velocity v;
        ev_index_query(&ix, tx[i].account, tx[i].ts, tx[i].seq_limit, &v);
        orc_features fv; float x[64]; igs_rule_row row; orc_score s;
        const igs_player_row* p = tx[i].account < n_players ? players + tx[i].account : NULL;
        orc_assemble(&tx[i], &v, p, &fv, x, &row);
        orc_score_features_row(c, m, x, &row, &s, NULL);
        to_result(&s, out + i);
        if (features_out) memcpy(features_out + i * 64, x, sizeof(x));
        if (velocity_out) velocity_out[i] = v;
    }
    ev_index_free(&ix);
}

/* ------------------------------------------------------------------------------------------ */
/* Indexed form of the same lookups — the timed CPU baseline of the fused workload.               */
/* ev_index_query above walks an account's events linearly (the obvious restatement; fine for     */
/* checking, hopeless on an account with 10^5 events).  Redis answers ZCOUNT in O(log n) from a   */
/* skiplist (features/redis_store.go:72-74), so a CPU baseline worth comparing against needs the   */
/* same complexity: per sorted event the running distinct count, per-type counts / sums, the      */
/* tx_sum chain and the session start are precomputed once, and a query is a few binary searches  */
/* and differences.  tests/ prove it equal to the linear closed form.                              */
/* ------------------------------------------------------------------------------------------ */
struct orc_fast_index {
    ev_index ix;
    uint32_t* dcnt;          /* [n+1] distinct (account, ts, amount) first occurrences before i      */
    uint32_t* cnt[2];        /* [n+1] bets / deposits before i                                        */
    int64_t*  sum[2];        /* [n+1]                                                                  */
    int64_t*  chain;         /* [n] tx_sum chain (gaps <= 3600) ending at i                            */
    uint32_t* sess;          /* [n] first ts of the gap <= 1800 chain ending at i                      */
};

typedef struct { int64_t amt; uint64_t pos; } amt_pos;
static int amt_pos_cmp(const void* a, const void* b) {
    const amt_pos* x = (const amt_pos*)a; const amt_pos* y = (const amt_pos*)b;
    if (x->amt != y->amt) return x->amt < y->amt ? -1 : 1;
    return x->pos < y->pos ? -1 : x->pos > y->pos;
}

orc_fast_index* orc_fast_index_build(const igs_event_cols* ev, uint64_t n, uint32_t n_accounts) {
    orc_fast_index* f = (orc_fast_index*)calloc(1, sizeof(*f));
    ev_index_build(&f->ix, ev, n, n_accounts);
    const ev_index* ix = &f->ix;
    f->dcnt = (uint32_t*)calloc(n + 1, 4);
    for (int t = 0; t < 2; ++t) {
        f->cnt[t] = (uint32_t*)calloc(n + 1, 4);
        f->sum[t] = (int64_t*)calloc(n + 1, 8);
    }
    f->chain = (int64_t*)calloc(n ? n : 1, 8);
    f->sess = (uint32_t*)calloc(n ? n : 1, 4);
    uint8_t* distinct = (uint8_t*)malloc(n ? n : 1);
    amt_pos* tmp = NULL; size_t tmp_cap = 0;
    for (uint64_t g0 = 0; g0 < n;) {                  /* (account, ts) groups: ZSET member dedupe */
        uint64_t g1 = g0 + 1;
        while (g1 < n && ix->keys[g1].account == ix->keys[g0].account && ix->ts[g1] == ix->ts[g0]) ++g1;
        const uint64_t len = g1 - g0;
        if (len <= 16) {
            for (uint64_t j = g0; j < g1; ++j) {
                distinct[j] = 1;
                for (uint64_t k = g0; k < j; ++k)
                    if (ix->amt[k] == ix->amt[j]) { distinct[j] = 0; break; }
            }
        } else {
            if (len > tmp_cap) { tmp_cap = len * 2; tmp = (amt_pos*)realloc(tmp, tmp_cap * sizeof(amt_pos)); }
            for (uint64_t j = 0; j < len; ++j) tmp[j] = (amt_pos){ix->amt[g0 + j], g0 + j};
            qsort(tmp, len, sizeof(amt_pos), amt_pos_cmp);
            for (uint64_t j = 0; j < len; ++j) distinct[tmp[j].pos] = j == 0 || tmp[j].amt != tmp[j - 1].amt;
        }
        g0 = g1;
    }
    free(tmp);
    for (uint64_t i = 0; i < n; ++i) {
        const int head = i == 0 || ix->keys[i].account != ix->keys[i - 1].account;
        const int64_t gap = head ? 0 : (int64_t)ix->ts[i] - (int64_t)ix->ts[i - 1];
        f->dcnt[i + 1] = f->dcnt[i] + distinct[i];
        for (int t = 0; t < 2; ++t) {
            f->cnt[t][i + 1] = f->cnt[t][i] + (ix->ty[i] == t);
            f->sum[t][i + 1] = f->sum[t][i] + (ix->ty[i] == t ? ix->amt[i] : 0);
        }
        f->chain[i] = (head || gap > 3600) ? ix->amt[i] : f->chain[i - 1] + ix->amt[i];
        f->sess[i] = (head || gap > 1800) ? ix->ts[i] : f->sess[i - 1];
    }
    free(distinct);
    return f;
}
void orc_fast_index_free(orc_fast_index* f) {
    if (!f) return;
    ev_index_free(&f->ix);
    free(f->dcnt); free(f->cnt[0]); free(f->cnt[1]); free(f->sum[0]); free(f->sum[1]); free(f->chain); free(f->sess);
    free(f);
}
static uint64_t fast_lower_ts(const uint32_t* ts, uint64_t lo, uint64_t hi, int64_t key) {   /* first ts >= key */
    while (lo < hi) {
        const uint64_t mid = lo + (hi - lo) / 2;
        if ((int64_t)ts[mid] < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}
void orc_fast_velocity(const orc_fast_index* f, uint32_t account, uint32_t ts_u, uint32_t seq_limit,
                       igs_velocity* v) {
    memset(v, 0, sizeof(*v));
    const ev_index* ix = &f->ix;
    if (account >= ix->n_accounts) return;
    const uint64_t s = ix->seg[account], e = ix->seg[account + 1];
    uint64_t lo = s, hi = e;                            /* prior events: (ts, seq) < (now, seq_limit) */
    while (lo < hi) {
        const uint64_t mid = lo + (hi - lo) / 2;
        if (ix->ts[mid] < ts_u || (ix->ts[mid] == ts_u && ix->seq[mid] < seq_limit)) lo = mid + 1; else hi = mid;
    }
    const uint64_t u = lo;
    v->n_prior = (uint32_t)(u - s);
    if (u == s) return;
    const int64_t now = ts_u, L = ix->ts[u - 1];
    const uint64_t alive = fast_lower_ts(ix->ts, s, u, L - 3599);        /* ZSET after the last trim: ts > L-3600 */
    const uint64_t b60 = fast_lower_ts(ix->ts, s, u, now - 60), b300 = fast_lower_ts(ix->ts, s, u, now - 300);
    const uint64_t b3600 = fast_lower_ts(ix->ts, s, u, now - 3600), b86400 = fast_lower_ts(ix->ts, s, u, now - 86400);
    #define ORC_MAX(a, b) ((a) > (b) ? (a) : (b))
    v->tx_count_1m = f->dcnt[u] - f->dcnt[ORC_MAX(b60, alive)];           /* ZCOUNT :72-74 */
    v->tx_count_5m = f->dcnt[u] - f->dcnt[ORC_MAX(b300, alive)];
    v->tx_count_1h = f->dcnt[u] - f->dcnt[ORC_MAX(b3600, alive)];
    #undef ORC_MAX
    const uint64_t wb[3] = {b60, b3600, b86400};
    for (int t = 0; t < 2; ++t)
        for (int w = 0; w < 3; ++w) {
            v->win[t][w].count = f->cnt[t][u] - f->cnt[t][wb[w]];
            v->win[t][w].sum = f->sum[t][u] - f->sum[t][wb[w]];
        }
    if (now - L <= 3600) v->tx_sum_1h = f->chain[u - 1];                  /* :77, :136-138 */
    if (now - L <= 7 * 24 * 3600) v->time_since_last_tx = (uint32_t)(now - L);
    if (now - L <= 1800) v->session_duration = (uint32_t)(now - (int64_t)f->sess[u - 1]);
}
void orc_fast_velocity_batch(const orc_fast_index* f, const igs_query* q, uint64_t n_q, igs_velocity* out) {
    for (uint64_t i = 0; i < n_q; ++i) orc_fast_velocity(f, q[i].account, q[i].ts, q[i].seq_limit, out + i);
}

typedef struct {
    const igs_config* c; const igs_gbdt_desc* m; const igs_player_row* players; uint64_t n_players;
    const orc_fast_index* f; const igs_tx* tx; uint64_t lo, hi; igs_result* out;
} sb_job;
static void* sb_worker(void* p) {
    sb_job* j = (sb_job*)p;
    for (uint64_t i = j->lo; i < j->hi; ++i) {           /* one request at a time: PredictBatch's loop, :314-326 */
        igs_velocity v;
        orc_fast_velocity(j->f, j->tx[i].account, j->tx[i].ts, j->tx[i].seq_limit, &v);
        orc_features fv; float x[64]; igs_rule_row row; orc_score s;
        const igs_player_row* pl = j->tx[i].account < j->n_players ? j->players + j->tx[i].account : NULL;
        orc_assemble(&j->tx[i], &v, pl, &fv, x, &row);
        orc_score_features_row(j->c, j->m, x, &row, &s, NULL);
        to_result(&s, j->out + i);
    }
    return NULL;
}
/* orc_score_batch against a prebuilt index, `threads` host threads over static chunks. */
void orc_score_batch_indexed(const igs_config* c, const igs_gbdt_desc* m, const igs_player_row* players,
                             uint64_t n_players, const orc_fast_index* f, const igs_tx* tx, uint64_t n,
                             igs_result* out, int threads) {
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t th[256];
    sb_job jobs[256];
    const uint64_t chunk = (n + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
        uint64_t lo = (uint64_t)t * chunk, hi = lo + chunk;
        if (lo > n) lo = n;
        if (hi > n) hi = n;
        jobs[t] = (sb_job){c, m, players, n_players, f, tx, lo, hi, out};
        if (threads == 1) sb_worker(&jobs[t]);
        else pthread_create(&th[t], NULL, sb_worker, &jobs[t]);
    }
    if (threads > 1)
        for (int t = 0; t < threads; ++t) pthread_join(th[t], NULL);
}

/* Path B with the identifier inputs computed here instead of arriving with the tx: uniques from
 * the history's identifier columns (when ev_ids != NULL), R8 from the blacklist sets (when
 * tx_ids != NULL; OR-ed with the tx's own IGS_BIT_BLACKLISTED). */
void orc_score_batch_ids(const igs_config* c, const igs_gbdt_desc* m, const igs_player_row* players,
                         uint64_t n_players, const igs_event_cols* ev, const igs_event_ids* ev_ids,
                         uint64_t n_events, const orc_blacklist* bl, const igs_tx* tx,
                         const igs_tx_ids* tx_ids, uint64_t n, igs_result* out, float* features_out) {
    ev_index ix;
    ev_index_build(&ix, ev, n_events, (uint32_t)n_players);
    if (ev_ids) ev_index_add_ids(&ix, ev_ids);
    for (uint64_t i = 0; i < n; ++i) {
        igs_tx t = tx[i];
        igs_velocity v;
        ev_index_query(&ix, t.account, t.ts, t.seq_limit, &v);
        if (ev_ids) {
            igs_uniques u;
            ev_index_uniques(&ix, t.account, t.ts, t.seq_limit, &u);
            t.unique_devices_24h = (uint16_t)(u.devices_24h > 65535 ? 65535 : u.devices_24h);
            t.unique_ips_24h = (uint16_t)(u.ips_24h > 65535 ? 65535 : u.ips_24h);
        }
        if (tx_ids && bl && orc_blacklist_check(bl, tx_ids + i)) t.bits |= IGS_BIT_BLACKLISTED;
        orc_features fv; float x[64]; igs_rule_row row; orc_score s;
        const igs_player_row* p = t.account < n_players ? players + t.account : NULL;
        orc_assemble(&t, &v, p, &fv, x, &row);
        orc_score_features_row(c, m, x, &row, &s, NULL);
        to_result(&s, out + i);
        if (features_out) memcpy(features_out + i * 64, x, sizeof(x));
    }
    ev_index_free(&ix);
}

/* ========================================================================================== */
/* linked accounts / bonus abuse.  The reference only declares the interfaces     */
/* (bonus_engine.go:139-141, risk.proto:135-145,274); the semantics are the spec in igs.h,        */
/* restated here the slow and obvious way: sort, unique, enumerate.  PARITY UNPINNED.             */
/* ========================================================================================== */
typedef struct { uint64_t a, b; } u64pair;
static int pair_cmp(const void* x, const void* y) {
    const u64pair* p = (const u64pair*)x; const u64pair* q = (const u64pair*)y;
    if (p->a != q->a) return p->a < q->a ? -1 : 1;
    if (p->b != q->b) return p->b < q->b ? -1 : 1;
    return 0;
}
static uint64_t pair_unique(u64pair* v, uint64_t n) {
    uint64_t w = 0;
    for (uint64_t i = 0; i < n; ++i)
        if (w == 0 || v[i].a != v[w - 1].a || v[i].b != v[w - 1].b) v[w++] = v[i];
    return w;
}

void orc_links(const uint64_t* device, const uint64_t* account, uint64_t n_events,
               const uint64_t* query, uint64_t n_q, igs_links_row* out) {
    u64pair* pr = (u64pair*)malloc((n_events ? n_events : 1) * sizeof(u64pair));
    uint64_t np = 0;
    for (uint64_t i = 0; i < n_events; ++i)
        if (device[i]) { pr[np].a = device[i]; pr[np].b = account[i]; np++; }   /* empty identifier: skipped */
    qsort(pr, np, sizeof(u64pair), pair_cmp);
    np = pair_unique(pr, np);
    uint64_t cap = 16, nl = 0;
    u64pair* lk = (u64pair*)malloc(cap * sizeof(u64pair));
    for (uint64_t s = 0; s < np;) {
        uint64_t e = s;
        while (e < np && pr[e].a == pr[s].a) ++e;
        const uint64_t k = e - s;                       /* accounts seen with this device */
        if (k >= 2 && k <= IGS_LINK_MAX_ACCOUNTS)
            for (uint64_t i = s; i < e; ++i)
                for (uint64_t j = s; j < e; ++j) {
                    if (i == j) continue;
                    if (nl == cap) { cap *= 2; lk = (u64pair*)realloc(lk, cap * sizeof(u64pair)); }
                    lk[nl].a = pr[i].b; lk[nl].b = pr[j].b; nl++;
                }
        s = e;
    }
    qsort(lk, nl, sizeof(u64pair), pair_cmp);
    nl = pair_unique(lk, nl);
    for (uint64_t q = 0; q < n_q; ++q) {
        igs_links_row r;
        memset(&r, 0, sizeof(r));
        uint64_t lo = 0, hi = nl;                       /* first link of query[q] */
        while (lo < hi) { uint64_t mid = (lo + hi) / 2; if (lk[mid].a < query[q]) lo = mid + 1; else hi = mid; }
        for (uint64_t i = lo; i < nl && lk[i].a == query[q]; ++i) {
            if (r.n_linked < 3) r.linked[r.n_linked] = lk[i].b;          /* sorted: the three smallest */
            r.n_linked++;
        }
        out[q] = r;
    }
    free(pr); free(lk);
}

void orc_check_bonus_abuse(const igs_player_row* p, const igs_links_row* l, igs_abuse* out) {
    memset(out, 0, sizeof(*out));
    const int bonus_only = p && p->bonus_claim_count > 3 && p->total_deposits < 5000;   /* engine.go:384-386 */
    const uint32_t nl = l ? l->n_linked : 0;
    uint32_t sig = 0;
    if (bonus_only) sig |= IGS_ABUSE_BONUS_ONLY;
    if (nl >= 1) sig |= IGS_ABUSE_MULTI_ACCOUNT;
    if (nl >= 3) sig |= IGS_ABUSE_ACCOUNT_RING;
    const int tenths = 5 * bonus_only + 3 * (nl >= 1) + 2 * (nl >= 3);
    out->abuse_score = (float)tenths / 10.0f;
    out->n_linked = nl;
    out->is_abuser = tenths >= 5;
    out->signals = (uint8_t)sig;
}

/* ========================================================================================== */
/* prediction/ltv.go                                                                           */
/* ========================================================================================== */

/* calculateEngagementScore, prediction/ltv.go:181-225. */
double orc_ltv_engagement(const igs_ltv_row* f) {
    double score = 0;
    if (f->days_since_last_bet < 3) score += 0.3;
    else if (f->days_since_last_bet < 7) score += 0.2;
    else if (f->days_since_last_bet < 14) score += 0.1;
    if (f->sessions_per_week >= 5) score += 0.2;
    else if (f->sessions_per_week >= 3) score += 0.15;
    else if (f->sessions_per_week >= 1) score += 0.1;
    if (f->deposit_frequency >= 4) score += 0.2;
    else if (f->deposit_frequency >= 2) score += 0.15;
    else if (f->deposit_frequency >= 1) score += 0.1;
    if (f->push_enabled) score += 0.1;
    if (f->email_opt_in) score += 0.1;
    if (f->has_vip_manager) score += 0.1;
    return score < 1.0 ? score : 1.0;                 /* math.Min(score, 1.0) */
}

/* calculateLTV, :155-178. */
double orc_ltv_heuristic(const igs_ltv_row* f) {
    double current = f->net_revenue;
    if (f->days_since_registration < 30) {
        int d = f->days_since_registration > 1 ? f->days_since_registration : 1;
        volatile double q = f->net_revenue / (double)d;
        volatile double monthly = q * 30;
        return monthly * 12;
    }
    volatile double q = f->net_revenue / (double)f->days_since_registration;
    volatile double monthly = q * 30;
    volatile double remaining = 12.0 * orc_ltv_engagement(f);
    volatile double future = monthly * remaining;
    return current + future;
}

/* calculateChurnRisk, :228-262. */
double orc_ltv_churn(const igs_ltv_row* f) {
    double risk = 0;
    if (f->days_since_last_bet > 30) risk += 0.5;
    else if (f->days_since_last_bet > 14) risk += 0.3;
    else if (f->days_since_last_bet > 7) risk += 0.15;
    if (f->sessions_per_week < 1 && f->days_since_registration > 30) risk += 0.2;
    if (f->days_since_last_deposit > 30) risk += 0.2;
    if (f->support_tickets > 3) risk += 0.1;
    if (f->total_withdrawals > f->total_deposits) risk += 0.1;
    return risk < 1.0 ? risk : 1.0;
}

/* Predict after the model output, :124-139; determineSegment :265-281; predictSurvival :284-297;
 * getNextBestAction :300-343; calculateConfidence :346-382. */
void orc_ltv_epilogue(const igs_ltv_row* f, double ltv, igs_ltv_result* out) {
    memset(out, 0, sizeof(*out));
    double churn = orc_ltv_churn(f);
    volatile double half = churn * 0.5;
    volatile double keep = 1 - half;
    double adjusted = ltv * keep;                                       /* :127 */
    int seg;
    if (churn > 0.7) seg = IGS_SEG_CHURNING;
    else if (adjusted >= 10000) seg = IGS_SEG_VIP;
    else if (adjusted >= 1000) seg = IGS_SEG_HIGH;
    else if (adjusted >= 100) seg = IGS_SEG_MEDIUM;
    else seg = IGS_SEG_LOW;
    volatile double em = 1.0 + orc_ltv_engagement(f);
    volatile double cm = 1.0 - churn;
    volatile double s1 = 90.0 * em;
    double surv = s1 * cm;
    if (!(surv > 0)) surv = 0;                                          /* math.Max(x, 0) */
    int act = 0;
    switch (seg) {
    case IGS_SEG_CHURNING: act = f->net_revenue > 0 ? 1 : 2; break;
    case IGS_SEG_VIP: act = f->days_since_last_deposit > 7 ? 3 : 4; break;
    case IGS_SEG_HIGH: act = !f->has_vip_manager ? 5 : (churn > 0.3 ? 6 : 7); break;
    case IGS_SEG_MEDIUM: act = f->bonuses_claimed < 3 ? 8 : (f->games_played < 5 ? 9 : 10); break;
    case IGS_SEG_LOW:
        act = f->days_since_registration < 7 ? 11 : (f->bonus_conversion_rate > 0.8 ? 0 : 12);
        break;
    }
    double conf = 0;
    if (f->days_since_registration > 90) conf += 0.3;
    else if (f->days_since_registration > 30) conf += 0.2;
    else conf += 0.1;
    if (f->bet_count > 100) conf += 0.3;
    else if (f->bet_count > 20) conf += 0.2;
    else conf += 0.1;
    if (f->deposit_frequency > 2) conf += 0.2;
    else if (f->deposit_frequency > 0) conf += 0.1;
    if (f->days_since_last_bet < 7) conf += 0.2;
    else if (f->days_since_last_bet < 30) conf += 0.1;
    if (conf > 1.0) conf = 1.0;
    out->predicted_ltv = adjusted;
    out->churn_risk = churn;
    out->confidence = (float)conf;                    /* wire type float, risk.proto:101 */
    out->days_survival = (int32_t)surv;
    out->segment = (uint8_t)seg;
    out->next_best_action = (uint8_t)act;
}

/* ------------------------------------------------------------------------------------------ */
/* ml.LTVModel (ml/onnx_model.go:366-490): the "model" the ml package returns for the LTV slot   */
/* while `session: nil // Use mock for now` (:399).  Its two helpers are NOT the predictor's      */
/* (prediction/ltv.go:181-262): no "< 14" / ">= 1" tiers, GamesPlayed >= 10 and                   */
/* BonusConversionRate > 0.5 instead of the notification flags, fewer churn terms.  The fields    */
/* are float32 and compare against untyped constants, i.e. in float32.                            */
/* ------------------------------------------------------------------------------------------ */
/* calculateEngagement, :434-467 */
double orc_ltv_model_engagement(const igs_ltv_features* f) {
    double score = 0;
    if (f->days_since_last_bet < 3) score += 0.3;                      /* :437-441 */
    else if (f->days_since_last_bet < 7) score += 0.2;
    if (f->sessions_per_week >= 5) score += 0.2;                       /* :443-447 */
    else if (f->sessions_per_week >= 3) score += 0.15;
    if (f->deposit_frequency >= 4) score += 0.2;                       /* :449-453 */
    else if (f->deposit_frequency >= 2) score += 0.15;
    if (f->games_played >= 10) score += 0.15;                          /* :455-457 */
    if (f->bonus_conversion_rate > 0.5f) score += 0.15;                /* :459-461 */
    if (score > 1) score = 1;                                          /* :463-465 */
    return score;
}
/* calculateChurnRisk, :469-490 */
double orc_ltv_model_churn(const igs_ltv_features* f) {
    double risk = 0;
    if (f->days_since_last_bet > 30) risk += 0.5;                      /* :472-476 */
    else if (f->days_since_last_bet > 14) risk += 0.3;
    if (f->days_since_last_deposit > 30) risk += 0.2;                  /* :478-480 */
    if (f->sessions_per_week < 1 && f->days_since_registration > 30) risk += 0.2;   /* :482-484 */
    if (risk > 1) risk = 1;                                            /* :486-488 */
    return risk;
}
/* Predict, :405-432 */
double orc_ltv_model_predict(const igs_ltv_features* f) {
    double current = (double)f->net_revenue;                           /* :412 */
    double days = (double)f->days_since_registration;                  /* :415-418 */
    if (days < 1) days = 1;
    volatile double q = current / days;
    volatile double monthly = q * 30;                                  /* :419 */
    volatile double remaining = 12.0 * orc_ltv_model_engagement(f);    /* :422-423 */
    volatile double half = orc_ltv_model_churn(f) * 0.5;
    volatile double adjust = 1.0 - half;                               /* :426-427 */
    volatile double future = monthly * remaining;
    volatile double total = current + future;
    return total * adjust;                                             /* :429 */
}

/* ========================================================================================== */
/* MLP spec (parity unpinned: stands in for ltv_model.onnx)                                    */
/* ========================================================================================== */
float orc_bf16_to_f32(uint16_t b) {
    union { uint32_t u; float f; } x; x.u = (uint32_t)b << 16; return x.f;
}
uint16_t orc_f32_to_bf16(float f) {
    union { uint32_t u; float f; } x; x.f = f;
    if ((x.u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((x.u >> 16) | 0x40);  /* NaN */
    uint32_t r = 0x7fffu + ((x.u >> 16) & 1u);
    return (uint16_t)((x.u + r) >> 16);
}

void orc_mlp_layer(const uint16_t* w, const float* b, int n_out, int n_in, const uint16_t* in,
                   float* pre) {
    for (int o = 0; o < n_out; ++o) {
        double acc = 0;
        const uint16_t* wr = w + (size_t)o * n_in;
        for (int i = 0; i < n_in; ++i) acc += (double)orc_bf16_to_f32(wr[i]) * (double)orc_bf16_to_f32(in[i]);
        volatile float s = (float)acc;
        pre[o] = s + b[o];
    }
}

float orc_mlp_forward(const igs_mlp_desc* m, const uint16_t* x, uint16_t* hidden_out) {
    static __thread float pre[IGS_MLP_HIDDEN];
    static __thread uint16_t h[2][IGS_MLP_HIDDEN];
    const uint16_t* W[3] = {m->w1, m->w2, m->w3};
    const float* B[3] = {m->b1, m->b2, m->b3};
    const uint16_t* in = x;
    int n_in = IGS_MLP_IN;
    for (int l = 0; l < 3; ++l) {
        orc_mlp_layer(W[l], B[l], IGS_MLP_HIDDEN, n_in, in, pre);
        uint16_t* o = h[l & 1];
        for (int i = 0; i < IGS_MLP_HIDDEN; ++i) o[i] = orc_f32_to_bf16(pre[i] > 0 ? pre[i] : 0.0f);
        if (hidden_out) memcpy(hidden_out + (size_t)l * IGS_MLP_HIDDEN, o, IGS_MLP_HIDDEN * 2);
        in = o; n_in = IGS_MLP_HIDDEN;
    }
    float y;
    orc_mlp_layer(m->w4, m->b4, 1, IGS_MLP_HIDDEN, in, &y);
    return y;
}
