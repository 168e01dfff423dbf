"""GPU: BASELINE.json's large configurations through size-independent properties + an oracle sample.

C4: LTV MLP over 4 000 000 players.  C5: one GPU's shard (12 500 000 tx, 500 000 players) of the 100M-tx
batch through the fused path with the 500-tree ensemble."""
import numpy as np
import pytest

from tests import synth

pytestmark = pytest.mark.gpu


def test_c4_ltv_four_million_players(engine, orc, pkg):
    n, period = 4_000_000, 65_536
    model = synth.make_mlp()
    engine.load_mlp(model)
    x, rows = synth.make_ltv_inputs(period)
    reps = -(-n // period)
    xb = np.tile(x, (reps, 1))[:n]
    rb = np.tile(rows, reps)[:n]
    out, raw = engine.ltv_batch(xb, rb, want_raw=True)
    # identical rows give bit-identical outputs wherever they sit (chunk, tile, epilogue warp)
    base = raw[:period].view(np.uint32)
    for k in range(1, reps):
        seg = raw[k * period:(k + 1) * period].view(np.uint32)
        assert np.array_equal(seg, base[:seg.shape[0]]), f"period {k} differs"
    assert out[:period].tobytes() == out[period:2 * period].tobytes()
    assert np.isfinite(raw).all()
    # oracle sample (tolerances as in tests/test_ltv_gpu.py)
    idx = np.arange(0, period, period // 64)[:64]
    want = orc.mlp_forward(model, x[idx])
    scale = np.maximum(np.maximum(1.0, np.abs(want)), np.sqrt(np.mean(want.astype(np.float64) ** 2)))
    err = np.abs(raw[idx].astype(np.float64) - want) / scale
    assert err.max() <= 1e-2 and np.percentile(err, 90) <= 1e-3
    # the epilogue is exact given the device's model output
    want_out = orc.ltv_epilogue(rows[idx], raw[idx].astype(np.float64))
    assert out[idx].tobytes() == want_out.tobytes()


def test_c5_one_shard_of_the_100m_batch(engine, orc, pkg):
    abi = pkg.abi
    n, n_players, n_events = 12_500_000, 500_000, 5_000_000
    players = synth.make_players(n_players, seed=71)
    ev = synth.make_events(n_events, n_players, seed=72)
    tx = synth.make_tx(n, n_players, seed=73, invalid_rate=0.001)
    model = synth.make_gbdt(n_trees=500)
    engine.load_players(players)
    engine.load_events(ev, n_players)
    engine.load_gbdt(model)
    cfg = pkg.default_config(model_kind=abi.MODEL_GBDT)
    res = engine.score_batch(cfg, tx)
    assert res.shape == (n,)
    assert ((res["action"] >= 1) & (res["action"] <= 3)).all() and (res["score"] <= 100).all()
    assert ((res["ml_score"] >= 0) & (res["ml_score"] <= 1)).all()
    # batch-split invariance and permutation equivariance: a result depends on its own tx only
    lo = engine.score_batch(cfg, tx[:1_000_003])
    assert lo.tobytes() == res[:1_000_003].tobytes()
    rng = np.random.default_rng(5)
    perm = rng.permutation(2_000_000)
    shuffled = engine.score_batch(cfg, np.ascontiguousarray(tx[:2_000_000][perm]))
    assert shuffled.tobytes() == res[:2_000_000][perm].tobytes()
    # same account + same inputs -> same result (checksum over duplicates of the first 1000 tx)
    dup = np.concatenate([tx[:1000], tx[:1000]])
    r2 = engine.score_batch(cfg, dup)
    assert r2[:1000].tobytes() == r2[1000:].tobytes() == res[:1000].tobytes()
    # oracle on a random sample (the oracle indexes the whole 5M-event history once)
    pick = rng.choice(n, size=4000, replace=False)
    want = orc.score_batch(cfg, model, players, ev, np.ascontiguousarray(tx[pick]))
    assert res[pick].tobytes() == want.tobytes()
    # the batch exercises the rules
    seen = int(np.bitwise_or.reduce(res["flags"][:2_000_000]))
    for bit in (abi.R_NEW_ACCOUNT_LARGE_TX, abi.R_MULTIPLE_DEVICES, abi.R_IP_COUNTRY_MISMATCH, abi.R_VPN_DETECTED,
                abi.R_KNOWN_FRAUDSTER, abi.R_BONUS_ABUSE):
        assert seen & bit
