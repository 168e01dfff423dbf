// ltv_math.cuh — the fp64 arithmetic of the LTV path, shared by the MLP epilogue (k_ltv.cu) and the
// model-free kernels (k_ltv_heur.cu).  Every operation is written out (no contraction: the library is
// built with --fmad=false and the multiplies / adds below are the _rn intrinsics), because Go on amd64
// does not fuse and the oracle is compiled with -ffp-contract=off.
//   prediction/ltv.go:155-382   LTVPredictor: calculateLTV, engagement, churn, segment, survival, action,
//                               confidence
//   ml/onnx_model.go:405-490    ml.LTVModel.Predict — the LTV counterpart of mockPredict ("model 0")
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "igs.h"

// calculateEngagementScore, prediction/ltv.go:181-225
static __device__ __forceinline__ double ltv_engagement(const igs_ltv_row& f) {
    double s = 0.0;
    if (f.days_since_last_bet < 3) s += 0.3; else if (f.days_since_last_bet < 7) s += 0.2;
    else if (f.days_since_last_bet < 14) s += 0.1;
    if (f.sessions_per_week >= 5) s += 0.2; else if (f.sessions_per_week >= 3) s += 0.15;
    else if (f.sessions_per_week >= 1) s += 0.1;
    if (f.deposit_frequency >= 4) s += 0.2; else if (f.deposit_frequency >= 2) s += 0.15;
    else if (f.deposit_frequency >= 1) s += 0.1;
    if (f.push_enabled) s += 0.1;
    if (f.email_opt_in) s += 0.1;
    if (f.has_vip_manager) s += 0.1;
    return s < 1.0 ? s : 1.0;
}
// calculateChurnRisk, :228-262
static __device__ __forceinline__ double ltv_churn(const igs_ltv_row& f) {
    double r = 0.0;
    if (f.days_since_last_bet > 30) r += 0.5; else if (f.days_since_last_bet > 14) r += 0.3;
    else if (f.days_since_last_bet > 7) r += 0.15;
    if (f.sessions_per_week < 1 && f.days_since_registration > 30) r += 0.2;
    if (f.days_since_last_deposit > 30) r += 0.2;
    if (f.support_tickets > 3) r += 0.1;
    if (f.total_withdrawals > f.total_deposits) r += 0.1;
    return r < 1.0 ? r : 1.0;
}
// calculateLTV, :155-178 — what the reference runs where the MLP would sit
static __device__ __forceinline__ double ltv_calculate(const igs_ltv_row& f) {
    if (f.days_since_registration < 30) {                                                   // :160-166
        const int d = f.days_since_registration > 1 ? f.days_since_registration : 1;
        const double monthly = __dmul_rn(__ddiv_rn(f.net_revenue, (double)d), 30.0);
        return __dmul_rn(monthly, 12.0);
    }
    const double monthly = __dmul_rn(__ddiv_rn(f.net_revenue, (double)f.days_since_registration), 30.0);   // :169
    const double remaining = __dmul_rn(12.0, ltv_engagement(f));                            // :172-173
    return __dadd_rn(f.net_revenue, __dmul_rn(monthly, remaining));                         // :175-177
}
// Predict after the model output, :124-139 (+ :265-382)
static __device__ __forceinline__ void ltv_epilogue(const igs_ltv_row& f, double ltv, igs_ltv_result* out) {
    const double churn = ltv_churn(f);
    const double adjusted = __dmul_rn(ltv, __dsub_rn(1.0, __dmul_rn(churn, 0.5)));          // :127
    int seg;
    if (churn > 0.7) seg = IGS_SEG_CHURNING;                                                // :267
    else if (adjusted >= 10000) seg = IGS_SEG_VIP;
    else if (adjusted >= 1000) seg = IGS_SEG_HIGH;
    else if (adjusted >= 100) seg = IGS_SEG_MEDIUM;
    else seg = IGS_SEG_LOW;
    double surv = __dmul_rn(__dmul_rn(90.0, __dadd_rn(1.0, ltv_engagement(f))), __dsub_rn(1.0, churn));  // :286-294
    if (!(surv > 0)) surv = 0;
    int act = 0;                                                                            // :300-343
    switch (seg) {
    case IGS_SEG_CHURNING: act = f.net_revenue > 0 ? 1 : 2; break;
    case IGS_SEG_VIP: act = f.days_since_last_deposit > 7 ? 3 : 4; break;
    case IGS_SEG_HIGH: act = !f.has_vip_manager ? 5 : (churn > 0.3 ? 6 : 7); break;
    case IGS_SEG_MEDIUM: act = f.bonuses_claimed < 3 ? 8 : (f.games_played < 5 ? 9 : 10); break;
    case IGS_SEG_LOW: act = f.days_since_registration < 7 ? 11 : (f.bonus_conversion_rate > 0.8 ? 0 : 12); break;
    }
    double conf = 0.0;                                                                      // :346-382
    if (f.days_since_registration > 90) conf += 0.3; else if (f.days_since_registration > 30) conf += 0.2;
    else conf += 0.1;
    if (f.bet_count > 100) conf += 0.3; else if (f.bet_count > 20) conf += 0.2; else conf += 0.1;
    if (f.deposit_frequency > 2) conf += 0.2; else if (f.deposit_frequency > 0) conf += 0.1;
    if (f.days_since_last_bet < 7) conf += 0.2; else if (f.days_since_last_bet < 30) conf += 0.1;
    if (conf > 1.0) conf = 1.0;
    out->predicted_ltv = adjusted;
    out->churn_risk = churn;
    out->confidence = (float)conf;
    out->days_survival = (int32_t)surv;
    out->segment = (uint8_t)seg;
    out->next_best_action = (uint8_t)act;
    for (int i = 0; i < 6; ++i) out->pad[i] = 0;
}

// ml.LTVModel.calculateEngagement, ml/onnx_model.go:434-467.  Not the predictor's score: no "< 14 days"
// tier, no ">= 1" tiers, and GamesPlayed / BonusConversionRate instead of the notification flags.
// The fields are float32 and compare against untyped constants, i.e. in float32 (0.5 is exact).
static __device__ __forceinline__ double ltv_mock_engagement(const igs_ltv_features& f) {
    double s = 0.0;
    if (f.days_since_last_bet < 3.f) s += 0.3; else if (f.days_since_last_bet < 7.f) s += 0.2;
    if (f.sessions_per_week >= 5.f) s += 0.2; else if (f.sessions_per_week >= 3.f) s += 0.15;
    if (f.deposit_frequency >= 4.f) s += 0.2; else if (f.deposit_frequency >= 2.f) s += 0.15;
    if (f.games_played >= 10.f) s += 0.15;
    if (f.bonus_conversion_rate > 0.5f) s += 0.15;
    return s > 1.0 ? 1.0 : s;
}
// ml.LTVModel.calculateChurnRisk, :469-490
static __device__ __forceinline__ double ltv_mock_churn(const igs_ltv_features& f) {
    double r = 0.0;
    if (f.days_since_last_bet > 30.f) r += 0.5; else if (f.days_since_last_bet > 14.f) r += 0.3;
    if (f.days_since_last_deposit > 30.f) r += 0.2;
    if (f.sessions_per_week < 1.f && f.days_since_registration > 30.f) r += 0.2;
    return r > 1.0 ? 1.0 : r;
}
// ml.LTVModel.Predict, :405-432
static __device__ __forceinline__ double ltv_mock_predict(const igs_ltv_features& f) {
    const double current = (double)f.net_revenue;                                           // :412
    double days = (double)f.days_since_registration;                                        // :415-418
    if (days < 1) days = 1;
    const double monthly = __dmul_rn(__ddiv_rn(current, days), 30.0);                       // :419
    const double remaining = __dmul_rn(12.0, ltv_mock_engagement(f));                       // :422-423
    const double adjust = __dsub_rn(1.0, __dmul_rn(ltv_mock_churn(f), 0.5));                // :426-427
    return __dmul_rn(__dadd_rn(current, __dmul_rn(monthly, remaining)), adjust);            // :429
}
