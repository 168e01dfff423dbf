"""CPU: pin the oracle against the hand-derived known-answer vectors of SURVEY.md §8c.

The reference ships no tests, fixtures or golden vectors, so tests/golden/kat_*.json
(derived by hand from the cited source lines) are the pins of the restatement.
"""
import ctypes as C
import json
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
TYPES = {"bet": 0, "deposit": 1, "withdraw": 2}
ACTIONS = {1: "approve", 2: "review", 3: "block"}
# emission order of applyRules (scoring/engine.go:424-475) then ML (:285-287)
EMISSION = [(0, "HIGH_VELOCITY"), (1, "NEW_ACCOUNT_LARGE_TX"), (3, "MULTIPLE_DEVICES"),
            (2, "IP_COUNTRY_MISMATCH"), (5, "VPN_DETECTED"), (7, "RAPID_DEPOSIT_WITHDRAW"),
            (8, "BONUS_ABUSE"), (6, "KNOWN_FRAUDSTER"), (9, "ML_HIGH_RISK")]


def reasons_of(flags: int):
    return [name for bit, name in EMISSION if flags & (1 << bit)]


def load(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def case_inputs(orc, pkg, case):
    """KAT case -> (cfg, model, x64, rule row) through the oracle's own mapping + Normalize."""
    abi = pkg.abi
    f = orc.Features()
    for k, v in case["features"].items():
        setattr(f, k, v)
    if f.tx_count_1h > 0:
        f.tx_avg_1h = f.tx_sum_1h / f.tx_count_1h
    ttype = TYPES[case["type"]]
    v30 = np.zeros(30, dtype=np.float32)
    orc.lib().orc_to_ml(C.byref(f), case["amount"], ttype, v30.ctypes.data)
    orc.lib().orc_normalize(v30.ctypes.data)
    x = np.zeros(64, dtype=np.float32)
    x[:30] = v30
    row = np.zeros(1, dtype=abi.rule_row_dtype)
    row["amount"] = case["amount"]
    row["total_deposits"] = f.total_deposits
    row["total_withdrawals"] = f.total_withdrawals
    row["tx_count_1m"] = f.tx_count_1m
    row["account_age_days"] = f.account_age
    row["time_since_last_tx"] = f.time_since_last_tx
    row["deposit_count"] = f.deposit_count
    row["unique_devices_24h"] = f.unique_devices_24h
    row["unique_ips_24h"] = f.unique_ips_24h
    row["tx_type"] = ttype
    row["bits"] = (abi.BIT_VALID | (abi.BIT_VPN if f.is_vpn else 0) | (abi.BIT_PROXY if f.is_proxy else 0)
                   | (abi.BIT_TOR if f.is_tor else 0) | (abi.BIT_BONUS_ONLY if f.bonus_only_player else 0)
                   | (abi.BIT_BLACKLISTED if case["blacklist"] else 0))
    model = None
    m = case["model"]
    if m == "nil":
        kind = abi.MODEL_NONE
    elif m == "mock":
        kind = abi.MODEL_MOCK
    else:
        kind = abi.MODEL_GBDT
        bits = int(m.split(":")[1], 16)
        val = np.array([bits], dtype=np.uint32).view(np.float32)[0]
        model = pkg.GbdtModel(np.zeros((1, 255), np.uint8), np.zeros((1, 255), np.float32),
                              np.full((1, 256), val, np.float32), base_score=0.0,
                              post_transform=abi.POST_NONE)
    cfg = orc.default_config(model_kind=kind)
    return cfg, model, x, row


@pytest.mark.parametrize("case", load("kat_scoring.json")["cases"], ids=lambda c: c["id"])
def test_kat_scoring(orc, pkg, case):
    cfg, model, x, row = case_inputs(orc, pkg, case)
    s = orc.score_row(cfg, model, x, row)
    assert s.rule_score == case["rule"]
    assert s.ml_score == pytest.approx(case["ml"], abs=1e-15)
    assert s.score == case["score"]
    assert ACTIONS[s.action] == case["action"]
    assert reasons_of(s.flags) == case["reasons"]


def test_kat_velocity_redis_emulation(orc):
    """Literal Redis-command emulation reproduces KAT set B row by row (score, then apply)."""
    kat = load("kat_velocity.json")
    L = orc.lib()
    acc = L.orc_redis_new()
    try:
        for r in kat["rows"]:
            now = kat["t0"] + r["dt"]
            rt = orc.Realtime()
            L.orc_redis_get(acc, now, C.byref(rt))
            tslt = now - rt.last_tx_timestamp if rt.last_tx_timestamp > 0 else 0
            sess = now - rt.session_start if rt.session_start > 0 else 0
            got = (rt.tx_count_1m, rt.tx_count_5m, rt.tx_count_1h, rt.tx_sum_1h, tslt, sess)
            want = (r["c1m"], r["c5m"], r["c1h"], r["sum1h"], r["tslt"], r["sess"])
            assert got == want, f"dt={r['dt']}"
            L.orc_redis_update(acc, now, r["amount"])
    finally:
        L.orc_redis_free(acc)


def test_kat_velocity_closed_form(orc, pkg):
    """The closed form the device kernel implements reproduces KAT set B through the batch API
    (history = all rows, query i sees the events that arrived before it: seq_limit = i)."""
    abi = pkg.abi
    kat = load("kat_velocity.json")
    rows = kat["rows"]
    n = len(rows)
    ev = pkg.Events(np.zeros(n, np.uint32), np.array([kat["t0"] + r["dt"] for r in rows], np.uint32),
                    np.array([r["amount"] for r in rows], np.int64), np.zeros(n, np.uint8))
    q = np.zeros(n, dtype=abi.query_dtype)
    q["ts"] = ev.ts
    q["seq_limit"] = np.arange(n)
    v = orc.velocity_batch(ev, 1, q)
    for i, r in enumerate(rows):
        got = tuple(int(v[i][k]) for k in ("tx_count_1m", "tx_count_5m", "tx_count_1h", "tx_sum_1h",
                                           "time_since_last_tx", "session_duration"))
        assert got == (r["c1m"], r["c5m"], r["c1h"], r["sum1h"], r["tslt"], r["sess"]), f"dt={r['dt']}"


def test_closed_form_equals_redis_emulation_on_random_streams(orc, pkg):
    """Property: on random single-account streams (bursts, equal seconds, equal amounts, gaps around
    every TTL) the closed form equals the command-level emulation — counts and sums bit-exact."""
    abi = pkg.abi
    L = orc.lib()
    rng = np.random.default_rng(12345)
    gaps_pool = np.array([0, 0, 0, 1, 2, 29, 30, 59, 60, 61, 299, 300, 301, 1799, 1800, 1801, 3599,
                          3600, 3601, 7199, 7200, 7201, 86399, 86400, 604799, 604800, 604801])
    for trial in range(60):
        n = int(rng.integers(1, 120))
        gaps = rng.choice(gaps_pool, size=n)
        if trial % 3 == 0:
            gaps = rng.integers(0, 90, size=n)
        ts = 1_700_000_000 + np.cumsum(gaps)
        amt = rng.choice(np.array([100, 200, 200, 300, 5000, -50]), size=n).astype(np.int64)
        ev = pkg.Events(np.zeros(n, np.uint32), ts.astype(np.uint32), amt, np.zeros(n, np.uint8))
        q = np.zeros(n, dtype=abi.query_dtype)
        q["ts"] = ev.ts
        q["seq_limit"] = np.arange(n)
        v = orc.velocity_batch(ev, 1, q)
        acc = L.orc_redis_new()
        try:
            for i in range(n):
                now = int(ts[i])
                rt = orc.Realtime()
                L.orc_redis_get(acc, now, C.byref(rt))
                tslt = now - rt.last_tx_timestamp if rt.last_tx_timestamp > 0 else 0
                sess = now - rt.session_start if rt.session_start > 0 else 0
                want = (rt.tx_count_1m, rt.tx_count_5m, rt.tx_count_1h, rt.tx_sum_1h, tslt, sess)
                got = tuple(int(v[i][k]) for k in ("tx_count_1m", "tx_count_5m", "tx_count_1h",
                                                   "tx_sum_1h", "time_since_last_tx", "session_duration"))
                assert got == want, f"trial {trial} event {i}"
                L.orc_redis_update(acc, now, int(amt[i]))
        finally:
            L.orc_redis_free(acc)


def test_query_between_events_equals_emulation(orc, pkg):
    """Pure queries (not applied) at arbitrary later times see the same state as Redis would."""
    abi = pkg.abi
    L = orc.lib()
    rng = np.random.default_rng(777)
    for trial in range(40):
        n = int(rng.integers(1, 40))
        ts = 1_700_000_000 + np.cumsum(rng.choice([0, 1, 45, 400, 1700, 1900, 3500, 3700], size=n))
        amt = rng.integers(1, 4, size=n).astype(np.int64) * 100
        ev = pkg.Events(np.zeros(n, np.uint32), ts.astype(np.uint32), amt, np.zeros(n, np.uint8))
        deltas = np.array([0, 1, 59, 60, 61, 300, 301, 1800, 1801, 3600, 3601, 7200, 7201, 604800, 604801])
        q = np.zeros(len(deltas), dtype=abi.query_dtype)
        q["ts"] = int(ts[-1]) + deltas
        q["seq_limit"] = abi.SEQ_ALL
        v = orc.velocity_batch(ev, 1, q)
        acc = L.orc_redis_new()
        try:
            for i in range(n):
                L.orc_redis_update(acc, int(ts[i]), int(amt[i]))
            # Redis expiry is lazy + monotone in wall clock: query in increasing time order
            for j, d in enumerate(deltas):
                now = int(ts[-1]) + int(d)
                rt = orc.Realtime()
                L.orc_redis_get(acc, now, C.byref(rt))
                tslt = now - rt.last_tx_timestamp if rt.last_tx_timestamp > 0 else 0
                sess = now - rt.session_start if rt.session_start > 0 else 0
                want = (rt.tx_count_1m, rt.tx_count_5m, rt.tx_count_1h, rt.tx_sum_1h, tslt, sess)
                got = tuple(int(v[j][k]) for k in ("tx_count_1m", "tx_count_5m", "tx_count_1h",
                                                   "tx_sum_1h", "time_since_last_tx", "session_duration"))
                assert got == want, f"trial {trial} delta {d}"
        finally:
            L.orc_redis_free(acc)


SEGS = {1: "vip", 2: "high", 3: "medium", 4: "low", 5: "churning"}
LTV_ACTIONS = ["NO_ACTION", "SEND_WINBACK_BONUS", "SEND_ENGAGEMENT_EMAIL", "VIP_MANAGER_CALL",
               "EXCLUSIVE_EVENT_INVITE", "ASSIGN_VIP_MANAGER", "RETENTION_BONUS", "LOYALTY_REWARD",
               "SUGGEST_BONUS", "RECOMMEND_NEW_GAMES", "STANDARD_PROMOTION", "ONBOARDING_GUIDE",
               "SMALL_DEPOSIT_BONUS"]


@pytest.mark.parametrize("case", load("kat_ltv.json")["cases"], ids=lambda c: c["id"])
def test_kat_ltv(orc, pkg, case):
    abi = pkg.abi
    row = np.zeros(1, dtype=abi.ltv_row_dtype)
    for k, v in case["f"].items():
        row[k] = v
    L = orc.lib()
    p = row.ctypes.data
    assert L.orc_ltv_engagement(p) == pytest.approx(case["engagement"], rel=1e-12)
    ltv = L.orc_ltv_heuristic(p)
    assert ltv == pytest.approx(case["ltv"], rel=1e-12)
    assert L.orc_ltv_churn(p) == pytest.approx(case["churn"], rel=1e-12, abs=1e-15)
    out = orc.ltv_epilogue(row, np.array([ltv]))[0]
    assert out["predicted_ltv"] == pytest.approx(case["adjusted"], rel=1e-12)
    assert SEGS[int(out["segment"])] == case["segment"]
    assert int(out["days_survival"]) == case["survival"]
    assert float(out["confidence"]) == pytest.approx(case["confidence"], rel=1e-6)
    assert LTV_ACTIONS[int(out["next_best_action"])] == case["action"]


@pytest.mark.parametrize("case", load("kat_ltv_model.json")["cases"], ids=lambda c: c["id"])
def test_kat_ltv_model(orc, pkg, case):
    """ml.LTVModel.Predict (ml/onnx_model.go:405-490) — exact: the vectors were derived by evaluating the cited
    lines left to right in float64."""
    row = np.zeros(1, dtype=pkg.abi.ltv_features_dtype)
    for k, v in case["f"].items():
        row[k] = v
    L = orc.lib()
    assert L.orc_ltv_model_engagement(row.ctypes.data) == case["engagement"]
    assert L.orc_ltv_model_churn(row.ctypes.data) == case["churn"]
    assert L.orc_ltv_model_predict(row.ctypes.data) == case["ltv"]
    assert orc.ltv_model_predict(row)[0] == case["ltv"]


def test_sigmoid_spec_is_accurate(orc):
    """The deterministic fp32 logistic stays within 2e-7 of the fp64 logistic on [-30, 30]."""
    xs = np.linspace(-30, 30, 20001).astype(np.float32)
    got = np.array([orc.sigmoid(x) for x in xs], dtype=np.float64)
    want = 1.0 / (1.0 + np.exp(-xs.astype(np.float64)))
    assert np.max(np.abs(got - want)) < 2e-7
    assert orc.sigmoid(0.0) == np.float32(0.5)
    assert orc.sigmoid(100.0) == np.float32(1.0)
    assert 0.0 <= float(orc.sigmoid(-100.0)) < 1e-37


def test_blend_is_not_contracted(orc, pkg):
    """K12/K13 are one float32 ulp apart and must land on 41 / 42 (engine.go:291-294 in fp64,
    multiply-then-add, truncation)."""
    cfg = orc.default_config(model_kind=pkg.abi.MODEL_GBDT)
    s = orc.Score()
    for bits, want in ((0x3F333333, 41), (0x3F333334, 42)):
        ml = float(np.array([bits], np.uint32).view(np.float32)[0])
        orc.lib().orc_blend(C.byref(cfg), 0, 0, ml, 1, C.byref(s))
        assert s.score == want


def test_account_age_matches_integer_days(orc, pkg):
    """int(time.Since(created).Hours()/24) == floor(seconds/86400) for every second offset near a
    day boundary (engine.go:374)."""
    abi = pkg.abi
    p = np.zeros(1, dtype=abi.player_row_dtype)
    f = orc.Features()
    now = 1_700_000_000
    for d in list(range(0, 5)) + [86399, 86400, 86401, 7 * 86400 - 1, 7 * 86400, 400 * 86400 + 17]:
        p["account_created_at"] = now - d
        orc.lib().orc_extract_features(None, p.ctypes.data, now, 0, C.byref(f))
        assert f.account_age == d // 86400


def test_indexed_cpu_port_equals_the_linear_closed_form(orc, pkg):
    """orc_fast_index (prefix columns + binary searches; the timed CPU baseline of the fused workload) against the
    linear restatement, on histories with forced (ts, amount) collisions, same-second bursts, every TTL boundary
    and stream-mode queries (seq_limit); then whole results through the multi-threaded scorer."""
    from tests import synth
    abi = pkg.abi
    rng = np.random.default_rng(12)
    n_acc, n = 300, 30000
    ev = synth.make_events(n, n_acc, seed=31, span=30 * 3600, collision_rate=0.05)
    ev.ts[:400] = ev.ts[0]                                   # one big same-second group (sort path of the dedupe)
    ev.account[:400] = 7
    ev.amount[:400] = rng.integers(1, 40, 400) * 100
    q = np.zeros(20000, dtype=abi.query_dtype)
    pick = rng.integers(0, n, q.shape[0])
    q["account"] = ev.account[pick]
    q["ts"] = ev.ts[pick] + rng.choice([0, 1, 59, 60, 61, 300, 1800, 1801, 3599, 3600, 3601, 86400, 86401, 700000],
                                       q.shape[0]).astype(np.uint32)
    q["seq_limit"] = abi.SEQ_ALL
    q["seq_limit"][::3] = rng.integers(0, n, q[::3].shape[0])
    q["account"][::501] = n_acc + 5                          # unknown account
    fast = orc.FastIndex(ev, n_acc)
    assert fast.velocity(q).tobytes() == orc.velocity_batch(ev, n_acc, q).tobytes()
    players = synth.make_players(n_acc, seed=32, now=synth.T0 + 30 * 3600)
    tx = synth.make_tx(5000, n_acc, seed=33, t_lo=synth.T0 + 3600, t_hi=synth.T0 + 30 * 3600, invalid_rate=0.02)
    model = synth.make_gbdt(n_trees=20, seed=34)
    cfg = orc.default_config(model_kind=abi.MODEL_GBDT)
    want = orc.score_batch(cfg, model, players, ev, tx)
    assert fast.score_batch(cfg, model, players, tx, threads=1).tobytes() == want.tobytes()
    assert fast.score_batch(cfg, model, players, tx, threads=5).tobytes() == want.tobytes()
    fast.close()
