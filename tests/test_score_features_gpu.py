"""GPU: fused rules + tree ensemble + blend kernel (path A) through the C ABI vs the oracle.

Bar (BASELINE.json north_star): rule flags and leaf indices bit-exact; float scores within 1e-5.
This implementation is stricter: the ml score is bit-identical too (shared deterministic fp32
logistic, leaves summed in tree order), hence the integer score and the action as well."""
import json
import os

import numpy as np
import pytest

from tests import synth
from tests.test_oracle_kat import ACTIONS, case_inputs, load, reasons_of

pytestmark = pytest.mark.gpu

ML_TOL = 1e-5   # north-star tolerance on the float score (we also assert bit equality)


def assert_same(got, want, got_leaves=None, want_leaves=None):
    assert np.array_equal(got["flags"], want["flags"]), "rule flags differ"
    if want_leaves is not None:
        assert np.array_equal(got_leaves, want_leaves), "leaf indices differ"
    assert np.max(np.abs(got["ml_score"].astype(np.float64) - want["ml_score"].astype(np.float64)),
                  initial=0.0) <= ML_TOL
    assert np.array_equal(got["ml_score"].view(np.uint32), want["ml_score"].view(np.uint32)), "ml bits differ"
    assert np.array_equal(got["score"], want["score"]), "scores differ"
    assert np.array_equal(got["action"], want["action"]), "actions differ"


def test_kat_set_a_on_device(engine, orc, pkg):
    """Every KAT of SURVEY §8c set A through igs_score_features."""
    for case in load("kat_scoring.json")["cases"]:
        cfg, model, x, row = case_inputs(orc, pkg, case)
        if model is not None:
            engine.load_gbdt(model)
        got = engine.score_features(cfg, x.reshape(1, 64), row)[0]
        lib = pkg.abi.load_library()
        import ctypes as C
        assert lib.igs_rule_score_from_flags(C.byref(cfg), int(got["flags"])) == case["rule"], case["id"]
        assert float(got["ml_score"]) == pytest.approx(case["ml"], abs=1e-7), case["id"]
        assert int(got["score"]) == case["score"], case["id"]
        assert ACTIONS[int(got["action"])] == case["action"], case["id"]
        assert reasons_of(int(got["flags"])) == case["reasons"], case["id"]


@pytest.mark.parametrize("n", [1, 31, 32, 255, 256, 257, 1000, 20000])
def test_gbdt_matches_oracle_ragged_sizes(engine, orc, pkg, n):
    model = synth.make_gbdt(n_trees=500, seed=3)
    feats, rows = synth.make_feature_batch(n, seed=100 + n, invalid_rate=0.02)
    cfg = pkg.default_config(model_kind=pkg.abi.MODEL_GBDT)
    engine.load_gbdt(model)
    got, leaves = engine.score_features(cfg, feats, rows, want_leaves=True)
    want, want_leaves = orc.score_features(cfg, model, feats, rows, want_leaves=True, threads=8)
    assert_same(got, want, leaves, want_leaves)


@pytest.mark.parametrize("n_trees", [0, 1, 4, 5, 6, 24, 25, 26, 49, 50, 51, 123])
def test_tree_count_tails(engine, orc, pkg, n_trees):
    """Chunk (25 trees) and ILP-group (5 trees) boundaries."""
    model = synth.make_gbdt(n_trees=n_trees, seed=50 + n_trees, base_score=0.25)
    feats = synth.make_unit_features(777, seed=n_trees)
    _, rows = synth.make_feature_batch(777, seed=9)
    cfg = pkg.default_config(model_kind=pkg.abi.MODEL_GBDT)
    engine.load_gbdt(model)
    got, leaves = engine.score_features(cfg, feats, rows, want_leaves=True)
    want, want_leaves = orc.score_features(cfg, model, feats, rows, want_leaves=True)
    assert_same(got, want, leaves, want_leaves)


def test_post_transform_none_and_extreme_margins(engine, orc, pkg):
    abi = pkg.abi
    for post in (abi.POST_NONE, abi.POST_LOGISTIC):
        model = synth.make_gbdt(n_trees=60, seed=77, post_transform=post)
        model.leaf = (model.leaf * 40).astype(np.float32)       # margins far outside [0,1] / saturating
        feats = synth.make_unit_features(5000, seed=4)
        _, rows = synth.make_feature_batch(5000, seed=5)
        cfg = pkg.default_config(model_kind=abi.MODEL_GBDT)
        engine.load_gbdt(model)
        got, leaves = engine.score_features(cfg, feats, rows, want_leaves=True)
        want, want_leaves = orc.score_features(cfg, model, feats, rows, want_leaves=True)
        assert_same(got, want, leaves, want_leaves)


def test_special_float_features(engine, orc, pkg):
    """NaN goes right, +-inf, +-0, denormals, thresholds equal to the feature value."""
    model = synth.make_gbdt(n_trees=100, seed=8)
    n = 4096
    feats = synth.make_unit_features(n, seed=6)
    rng = np.random.default_rng(5)
    specials = np.array([np.nan, np.inf, -np.inf, 0.0, -0.0, 1e-45, -1e-45, 1.0, 3.4e38], dtype=np.float32)
    mask = rng.random(feats.shape) < 0.2
    feats[mask] = rng.choice(specials, size=int(mask.sum()))
    # plant exact threshold hits: x == thr must go left
    for i in range(0, n, 7):
        t, node = rng.integers(0, 100), rng.integers(0, 255)
        feats[i, model.feat[t, node]] = model.thr[t, node]
    _, rows = synth.make_feature_batch(n, seed=7)
    cfg = pkg.default_config(model_kind=pkg.abi.MODEL_GBDT)
    engine.load_gbdt(model)
    got, leaves = engine.score_features(cfg, feats, rows, want_leaves=True)
    want, want_leaves = orc.score_features(cfg, model, feats, rows, want_leaves=True)
    assert_same(got, want, leaves, want_leaves)


@pytest.mark.parametrize("kind", ["none", "mock"])
def test_rules_and_mock_model(engine, orc, pkg, kind):
    abi = pkg.abi
    n = 50000
    feats, rows = synth.make_feature_batch(n, seed=21, invalid_rate=0.05)
    cfg = pkg.default_config(model_kind=abi.MODEL_NONE if kind == "none" else abi.MODEL_MOCK)
    got = engine.score_features(cfg, feats, rows)
    want = orc.score_features(cfg, None, feats, rows, threads=8)
    assert_same(got, want)
    # the synthetic batch must actually exercise every rule the reference can emit
    seen = np.bitwise_or.reduce(want["flags"])
    for bit in (abi.R_HIGH_VELOCITY, abi.R_NEW_ACCOUNT_LARGE_TX, abi.R_IP_COUNTRY_MISMATCH,
                abi.R_MULTIPLE_DEVICES, abi.R_VPN_DETECTED, abi.R_KNOWN_FRAUDSTER,
                abi.R_RAPID_DEPOSIT_WITHDRAW, abi.R_BONUS_ABUSE):
        assert seen & bit, f"rule bit {bit} never fired in the synthetic batch"
    assert not (seen & abi.R_SUSPICIOUS_PATTERN)
    if kind == "mock":
        assert seen & abi.R_ML_HIGH_RISK


def test_config_overrides(engine, orc, pkg):
    """Thresholds / weights are per-call values (SetThresholds, engine.go:498-504)."""
    abi = pkg.abi
    model = synth.make_gbdt(n_trees=50, seed=31)
    feats, rows = synth.make_feature_batch(8192, seed=32)
    engine.load_gbdt(model)
    for kw in (dict(block_threshold=30, review_threshold=10),
               dict(ml_weight=0.9, rule_weight=0.1),
               dict(max_tx_per_minute=2, new_account_days=400, large_deposit_amount=500,
                    max_devices_per_day=1, max_ips_per_day=1),
               dict(reason_weight=[5, 6, 7, 8, 9, 10, 11, 12, 13, 14])):
        cfg = pkg.default_config(model_kind=abi.MODEL_GBDT, **kw)
        ocfg = orc.default_config(model_kind=abi.MODEL_GBDT, **kw)
        got = engine.score_features(cfg, feats, rows)
        want = orc.score_features(ocfg, model, feats, rows, threads=8)
        assert_same(got, want)


def test_empty_batch_and_errors(engine, pkg):
    abi = pkg.abi
    cfg = pkg.default_config(model_kind=abi.MODEL_MOCK)
    out = engine.score_features(cfg, np.zeros((0, 64), np.float32), np.zeros(0, abi.rule_row_dtype))
    assert out.shape == (0,)
    with pkg.Engine(device=0) as fresh:
        gcfg = pkg.default_config(model_kind=abi.MODEL_GBDT)
        with pytest.raises(pkg.IgsError) as ei:
            fresh.score_features(gcfg, np.zeros((4, 64), np.float32), np.zeros(4, abi.rule_row_dtype))
        assert ei.value.status == abi.ERR_STATE
        bad = synth.make_gbdt(n_trees=2, seed=1)
        bad.leaf[1, 3] = np.nan
        with pytest.raises(pkg.IgsError) as ei:
            fresh.load_gbdt(bad)
        assert ei.value.status == abi.ERR_INVALID


def test_full_size_c2_against_oracle_and_determinism(engine, orc, pkg):
    """BASELINE config 2: 1M tx x 64 features, 500 trees depth 8.  The oracle on all host cores
    finishes this in seconds, so the full batch is compared, not a sample."""
    n = 1_000_000
    model = synth.make_gbdt(n_trees=500)
    feats, rows = synth.make_feature_batch(n)
    cfg = pkg.default_config(model_kind=pkg.abi.MODEL_GBDT)
    engine.load_gbdt(model)
    got = engine.score_features(cfg, feats, rows)
    again = engine.score_features(cfg, feats, rows)
    assert got.tobytes() == again.tobytes(), "two runs of the same batch differ"
    want = orc.score_features(cfg, model, feats, rows, threads=os.cpu_count() or 8)
    assert_same(got, want)
    # checksum of checksums: order-independent digest of the result column bytes
    assert int(got.view(np.uint64).sum(dtype=np.uint64)) == int(want.view(np.uint64).sum(dtype=np.uint64))
