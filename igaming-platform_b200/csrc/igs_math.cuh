// igs_math.cuh — device-side arithmetic of the scoring path.  Every function names the reference
// lines it reproduces (paths relative to the reference root).  The library is compiled with
// --fmad=false; where a fused multiply-add is wanted it is written explicitly, so the float and
// double sequences below are exactly the IEEE operations listed, in that order.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "igs.h"

// Rule evaluation: applyRules, scoring/engine.go:420-483.  All comparisons strict; weights are
// indexed by reason bit.  Returns the capped rule score, *flags gets the rule bits.
__device__ __forceinline__ int igs_apply_rules(const igs_config& c, long long amount, int tx_type,
                                               unsigned tx_count_1m, int account_age,
                                               unsigned uniq_dev, unsigned uniq_ip, unsigned bits,
                                               unsigned tslt, unsigned deposit_count,
                                               long long total_deposits, long long total_withdrawals,
                                               unsigned* flags) {
    int total = 0;
    unsigned fl = 0;
    if ((long long)tx_count_1m > (long long)c.max_tx_per_minute) {                 // R1 :425
        total += c.reason_weight[0]; fl |= IGS_R_HIGH_VELOCITY;
    }
    if (account_age < c.new_account_days && amount > c.large_deposit_amount) {     // R2 :431
        total += c.reason_weight[1]; fl |= IGS_R_NEW_ACCOUNT_LARGE_TX;
    }
    if ((int)uniq_dev > c.max_devices_per_day) {                                   // R3 :437
        total += c.reason_weight[3]; fl |= IGS_R_MULTIPLE_DEVICES;
    }
    if ((int)uniq_ip > c.max_ips_per_day) {                                        // R4 :443
        total += c.reason_weight[2]; fl |= IGS_R_IP_COUNTRY_MISMATCH;
    }
    if (bits & (IGS_BIT_VPN | IGS_BIT_PROXY | IGS_BIT_TOR)) {                      // R5 :449
        total += c.reason_weight[5]; fl |= IGS_R_VPN_DETECTED;
    }
    if (tslt < 300u && tx_type == IGS_TX_WITHDRAW) {                               // R6 :455-460
        // int64 multiply, then truncating divide (Go wraps on overflow; so does this)
        long long lim = (long long)((unsigned long long)total_deposits * 80ull) / 100;
        if (deposit_count > 0 && total_withdrawals > lim) {
            total += c.reason_weight[7]; fl |= IGS_R_RAPID_DEPOSIT_WITHDRAW;
        }
    }
    if (bits & IGS_BIT_BONUS_ONLY) {                                               // R7 :463
        total += c.reason_weight[8]; fl |= IGS_R_BONUS_ABUSE;
    }
    if (bits & IGS_BIT_BLACKLISTED) {                                              // R8 :469-475
        total += c.reason_weight[6]; fl |= IGS_R_KNOWN_FRAUDSTER;
    }
    if (total > 100) total = 100;                                                  // :478
    *flags = fl;
    return total;
}

// minMaxScale(x, 0, hi), ml/onnx_model.go:197-205 (IEEE divide).
__device__ __forceinline__ float igs_min_max(float x, float hi) {
    if (x < 0.f) return 0.f;
    if (x > hi) return 1.f;
    return __fdiv_rn(x, hi);
}
// logTransform, ml/onnx_model.go:186-195 — log1p there is the identity.
__device__ __forceinline__ float igs_log_transform(float x) { return x <= 0.f ? 0.f : x; }

// Deterministic fp32 logistic — the spec shared with the oracle (oracle/igs_oracle.c
// orc_sigmoid): one correctly-rounded IEEE op per step, so CPU and GPU agree bit for bit.
__device__ __forceinline__ float igs_sigmoid(float m) {
    float z = -m;
    if (!(z >= -87.0f)) z = -87.0f;
    if (z > 88.0f) z = 88.0f;
    float t = __fmul_rn(z, 1.44269504088896341f);
    float n = rintf(t);
    float r = __fmaf_rn(n, -0.693145751953125f, z);
    r = __fmaf_rn(n, -1.42860682030941723e-06f, r);
    float p = 1.0f / 5040.0f;
    p = __fmaf_rn(p, r, 1.0f / 720.0f);
    p = __fmaf_rn(p, r, 1.0f / 120.0f);
    p = __fmaf_rn(p, r, 1.0f / 24.0f);
    p = __fmaf_rn(p, r, 1.0f / 6.0f);
    p = __fmaf_rn(p, r, 0.5f);
    p = __fmaf_rn(p, r, 1.0f);
    p = __fmaf_rn(p, r, 1.0f);
    float sc = __int_as_float(((int)n + 127) << 23);
    float e = __fmul_rn(p, sc);
    float d = __fadd_rn(1.0f, e);
    return __fdiv_rn(1.0f, d);
}

// Blend + action, scoring/engine.go:284-310.  fp64, multiply-then-add, truncation.
__device__ __forceinline__ void igs_blend(const igs_config& c, int rule_score, unsigned rule_flags,
                                          double ml, bool has_model, igs_result* out) {
    unsigned fl = rule_flags;
    if (!has_model) ml = 0.0;
    else if (ml > 0.7) fl |= IGS_R_ML_HIGH_RISK;                                   // :285-287
    double a = __dmul_rn(c.rule_weight, (double)rule_score);                       // :291-294
    double b = __dmul_rn(c.ml_weight, __dmul_rn(ml, 100.0));
    double s = __dadd_rn(a, b);
    long long fs = __double2ll_rz(s);
    if (fs > 100) fs = 100;                                                        // :297-299
    // The reference would return a negative int for a negative blend (negative weights): below every
    // threshold, i.e. approve.  The result byte cannot hold it; the action is taken first, then the score is
    // clamped at 0.
    int action = fs >= c.block_threshold ? IGS_ACTION_BLOCK
               : fs >= c.review_threshold ? IGS_ACTION_REVIEW : IGS_ACTION_APPROVE; // :303-310
    out->ml_score = (float)ml;
    out->flags = (uint16_t)fl;
    out->score = (uint8_t)(fs < 0 ? 0 : fs);
    out->action = (uint8_t)action;
}
