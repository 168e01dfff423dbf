"""GPU: incremental history ingest (igs_append_events) — the batch form of ScoringEngine.UpdateFeatures
(scoring/engine.go:486-488 -> features/redis_store.go:119-168).  Appending must be indistinguishable from
loading the concatenated history, and a score-then-append stream must match the oracle step by step."""
import numpy as np
import pytest

from tests import synth

pytestmark = pytest.mark.gpu


def cut(ev, lo, hi, pkg):
    return pkg.Events(ev.account[lo:hi], ev.ts[lo:hi], ev.amount[lo:hi], ev.tx_type[lo:hi])


def test_append_equals_bulk_load(engine, orc, pkg):
    abi = pkg.abi
    n_acc, n = 700, 30000
    ev = synth.make_events(n, n_acc, seed=41, span=5 * 3600, collision_rate=0.04)
    rng = np.random.default_rng(4)
    q = np.zeros(6000, dtype=abi.query_dtype)
    pick = rng.integers(0, n, size=q.shape[0])
    q["account"] = ev.account[pick]
    q["ts"] = ev.ts[pick] + rng.choice([0, 1, 60, 61, 3600, 3601, 90000], size=q.shape[0]).astype(np.uint32)
    q["seq_limit"] = abi.SEQ_ALL
    q["seq_limit"][::3] = rng.integers(0, n, size=q[::3].shape[0])
    engine.load_events(ev, n_acc)
    bulk = engine.get_velocity(q)
    # same history in uneven pieces, starting from an empty load
    engine.load_events(cut(ev, 0, 0, pkg), n_acc)
    for lo, hi in ((0, 1), (1, 2049), (2049, 2049), (2049, 17000), (17000, n)):
        engine.append_events(cut(ev, lo, hi, pkg))
    assert engine.get_velocity(q).tobytes() == bulk.tobytes()
    assert bulk.tobytes() == orc.velocity_batch(ev, n_acc, q).tobytes()


def test_score_then_append_stream_matches_oracle(engine, orc, pkg):
    """Micro-batches in time order: score batch k against the history of batches < k, then append it."""
    abi = pkg.abi
    n_acc, n, step = 150, 8000, 1000
    players = synth.make_players(n_acc, seed=51, now=synth.T0 + 2 * 3600)
    ev = synth.make_events(n, n_acc, seed=52, span=2 * 3600, zipf_s=0.7)
    order = np.argsort(ev.ts, kind="stable")
    ev = pkg.Events(ev.account[order], ev.ts[order], ev.amount[order], ev.tx_type[order])
    engine.load_players(players)
    engine.load_events(cut(ev, 0, 0, pkg), n_acc)
    cfg = pkg.default_config(model_kind=abi.MODEL_MOCK)
    fired = 0
    for lo in range(0, n, step):
        hi = lo + step
        tx = np.zeros(step, dtype=abi.tx_dtype)
        tx["account"], tx["ts"] = ev.account[lo:hi], ev.ts[lo:hi]
        tx["amount"], tx["tx_type"] = ev.amount[lo:hi], ev.tx_type[lo:hi]
        tx["seq_limit"] = abi.SEQ_ALL
        tx["unique_devices_24h"] = 1
        tx["unique_ips_24h"] = 1
        tx["bits"] = abi.BIT_VALID
        got = engine.score_batch(cfg, tx)
        want = orc.score_batch(cfg, None, players, cut(ev, 0, lo, pkg), tx)
        assert got.tobytes() == want.tobytes(), f"batch starting at {lo}"
        fired += int((want["flags"] & abi.R_HIGH_VELOCITY).astype(bool).sum())
        engine.append_events(cut(ev, lo, hi, pkg))
    assert fired > 0


def test_append_errors(engine, pkg):
    abi = pkg.abi
    with pkg.Engine(device=0) as fresh:
        one = pkg.Events(np.array([0], np.uint32), np.array([5], np.uint32), np.array([1], np.int64), np.zeros(1, np.uint8))
        with pytest.raises(pkg.IgsError) as ei:
            fresh.append_events(one)
        assert ei.value.status == abi.ERR_STATE
        fresh.load_events(one, 3)
        bad = pkg.Events(np.array([1, 7], np.uint32), np.array([6, 7], np.uint32), np.array([1, 1], np.int64), np.zeros(2, np.uint8))
        with pytest.raises(pkg.IgsError) as ei:
            fresh.append_events(bad)
        assert ei.value.status == abi.ERR_INVALID
        q = np.zeros(3, dtype=abi.query_dtype)
        q["account"] = [0, 1, 2]
        q["ts"] = 100
        q["seq_limit"] = abi.SEQ_ALL
        assert fresh.get_velocity(q)["n_prior"].tolist() == [1, 0, 0]      # the rejected batch left no trace
        fresh.append_events(pkg.Events(np.array([1], np.uint32), np.array([6], np.uint32), np.array([1], np.int64), np.zeros(1, np.uint8)))
        assert fresh.get_velocity(q)["n_prior"].tolist() == [1, 1, 0]
