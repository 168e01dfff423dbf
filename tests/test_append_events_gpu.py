"""GPU: incremental history ingest (igs_append_events) — the batch form of ScoringEngine.UpdateFeatures
(scoring/engine.go:486-488 -> features/redis_store.go:119-168).  Appending must be indistinguishable from
loading the concatenated history, and a score-then-append stream must match the oracle step by step."""
import os

import numpy as np
import pytest

from tests import synth

pytestmark = pytest.mark.gpu


def cut(ev, lo, hi, pkg):
    return pkg.Events(ev.account[lo:hi], ev.ts[lo:hi], ev.amount[lo:hi], ev.tx_type[lo:hi])


def test_append_equals_bulk_load(engine, orc, pkg, late_reorder):
    """Pieces of a history in random time order: with IGS_LATE_REORDER an append is a bulk load of the whole."""
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


def test_late_events_policy(engine, orc, pkg):
    """A late event — older than an event of its account that arrived before it — is where the index (events
    filed by timestamp) and Redis (events applied on arrival, features/redis_store.go:121-160) part ways.
    Smallest case, one account: arrivals (T0+100, 500) then (T0+10, 200); query at T0+120.  Applied on arrival,
    the second event overwrites last_tx with the OLDER time (:155): time_since_last_tx = 110.  Filed by timestamp:
    20.  igs_append_events refuses such a batch unless IGS_LATE_REORDER is set; igs_load_events (a replayed log)
    orders by timestamp and reports the late pair."""
    abi = pkg.abi
    T0 = synth.T0
    # the deviation itself, on the command-level emulation (wall clock == event timestamp, as everywhere)
    acc = orc.lib().orc_redis_new()
    rt = orc.Realtime()
    import ctypes as C
    for ts, amt in ((T0 + 100, 500), (T0 + 10, 200)):
        orc.lib().orc_redis_update(acc, ts, amt)
    orc.lib().orc_redis_get(acc, T0 + 120, C.byref(rt))
    orc.lib().orc_redis_free(acc)
    assert T0 + 120 - rt.last_tx_timestamp == 110 and rt.tx_count_5m == 2 and rt.tx_sum_1h == 700

    def one(ts, amt):
        return pkg.Events(np.array([0], np.uint32), np.array([ts], np.uint32), np.array([amt], np.int64), np.zeros(1, np.uint8))
    q = np.zeros(1, dtype=abi.query_dtype)
    q["ts"], q["seq_limit"] = T0 + 120, abi.SEQ_ALL
    engine.load_events(one(T0 + 100, 500), 1)
    with pytest.raises(pkg.IgsError) as ei:
        engine.append_events(one(T0 + 10, 200))
    assert ei.value.status == abi.ERR_INVALID and "late" in str(ei.value)
    v = engine.get_velocity(q)[0]
    assert (int(v["n_prior"]), int(v["time_since_last_tx"])) == (1, 20)          # the rejected event left no trace
    assert engine.history_info()["late_events"] == 0
    # late inside one batch: (T0+130) then (T0+125) of the same account
    two = pkg.Events(np.zeros(2, np.uint32), np.array([T0 + 130, T0 + 125], np.uint32), np.array([1, 2], np.int64), np.zeros(2, np.uint8))
    with pytest.raises(pkg.IgsError):
        engine.append_events(two)
    # equal timestamps are not late; neither is an old event of ANOTHER account
    engine.load_events(one(T0 + 100, 500), 2)
    engine.append_events(one(T0 + 100, 300))
    engine.append_events(pkg.Events(np.array([1], np.uint32), np.array([T0 + 5], np.uint32), np.array([9], np.int64), np.zeros(1, np.uint8)))
    info = engine.history_info()
    assert info.pop("delta_events") in (2, 0)             # 0 only under the IGS_SORT_PAIRS test knob (no delta index)
    assert info == {"n_events": 3, "late_events": 0, "n_accounts": 2, "has_identifiers": False}
    # opt in: accepted and filed by timestamp == bulk load == the closed form on the time-ordered history
    engine.set_late_policy(abi.LATE_REORDER)
    try:
        engine.load_events(one(T0 + 100, 500), 1)
        engine.append_events(one(T0 + 10, 200))
        got = engine.get_velocity(q)
        assert engine.history_info()["late_events"] == 1
    finally:
        engine.set_late_policy(abi.LATE_REJECT)
    both = pkg.Events(np.zeros(2, np.uint32), np.array([T0 + 100, T0 + 10], np.uint32), np.array([500, 200], np.int64), np.zeros(2, np.uint8))
    assert got.tobytes() == orc.velocity_batch(both, 1, q).tobytes()
    assert int(got[0]["time_since_last_tx"]) == 20 and int(got[0]["tx_count_5m"]) == 2
    engine.load_events(both, 1)                                # bulk load: accepted under either policy, counted
    assert engine.history_info()["late_events"] == 1
    assert engine.get_velocity(q).tobytes() == got.tobytes()


def test_base_delta_boundary_kats(engine, orc, pkg):
    """Appended events live in a delta index next to the base (igs_append_events): hand-built cases where a
    quantity straddles the boundary, each checked against the oracle on the concatenated history and against a
    bulk load of it.  One account per case; base = what igs_load_events gets, delta = what is appended."""
    abi = pkg.abi
    T = synth.T0 + 10_000
    cases = {
        # Q1: the ZSET member "ts:amount" exists in the base; its duplicates in the delta must not count again
        "q1_duplicate_split": ([(T, 100), (T, 200)], [(T, 100), (T, 300), (T, 200), (T, 100)]),
        # chain sum (tx_sum:1h, gaps <= 3600) and session (gaps <= 1800) continue across the boundary ...
        "chains_continue": ([(T - 5000, 7), (T - 1800, 11)], [(T, 13), (T + 1800, 17)]),
        # ... and break exactly one second later
        "chains_break": ([(T - 5000, 7), (T - 3601, 11)], [(T, 13), (T + 1801, 17)]),
        "chain_at_3600": ([(T - 3600, 11)], [(T, 13)]),
        # an event exactly at t - 3600 / t - 60 / t - 300 with the window bound falling on the boundary
        "window_edges": ([(T - 3600, 1), (T - 300, 2), (T - 60, 3)], [(T - 60, 4), (T, 5)]),
        # Q4: the newest prior event has ts == now, so the 1h trim bound (now - 3599) beats the window bound
        "q4_trim": ([(T - 3600, 1), (T - 3599, 2)], [(T, 3)]),
        # everything in the delta, nothing in the base, and the reverse
        "delta_only": ([], [(T - 100, 1), (T - 50, 2), (T, 3)]),
        "base_only": ([(T - 100, 1), (T - 50, 2), (T, 3)], []),
        # windows that end inside the base while the prior events reach the delta
        "long_tail": ([(T - 90000, 1), (T - 86400, 2), (T - 86399, 3), (T - 4000, 4)], [(T - 10, 5)]),
    }
    names = list(cases)
    base_rows = [(a, ts, amt) for a, k in enumerate(names) for ts, amt in cases[k][0]]
    delta_rows = [(a, ts, amt) for a, k in enumerate(names) for ts, amt in cases[k][1]]
    delta_rows.sort(key=lambda r: r[1])                      # arrival in time order

    def events(rows):
        a = np.array([r[0] for r in rows], np.uint32)
        return pkg.Events(a, np.array([r[1] for r in rows], np.uint32), np.array([r[2] for r in rows], np.int64),
                          (np.arange(len(rows)) % 3).astype(np.uint8))
    base, delta = events(base_rows), events(delta_rows)
    n_acc = len(names)
    q_ts = [T - 1, T, T + 1, T + 59, T + 60, T + 61, T + 300, T + 1800, T + 1801, T + 3599, T + 3600, T + 3601,
            T + 86399, T + 86400, T + 86401, T + 8 * 86400]
    q = np.zeros(n_acc * len(q_ts) * 2, dtype=abi.query_dtype)
    q["account"] = np.repeat(np.arange(n_acc), len(q_ts) * 2)
    q["ts"] = np.tile(np.repeat(q_ts, 2), n_acc)
    q["seq_limit"] = abi.SEQ_ALL
    q["seq_limit"][1::2] = np.tile(np.arange(len(q_ts)) % (len(base) + len(delta) + 1), n_acc)   # stream-mode cuts
    full = pkg.Events(np.concatenate([base.account, delta.account]), np.concatenate([base.ts, delta.ts]),
                      np.concatenate([base.amount, delta.amount]), np.concatenate([base.tx_type, delta.tx_type]))
    want = orc.velocity_batch(full, n_acc, q)
    engine.load_events(base, n_acc)
    half = len(delta) // 2
    engine.append_events(pkg.Events(delta.account[:half], delta.ts[:half], delta.amount[:half], delta.tx_type[:half]))
    engine.append_events(pkg.Events(delta.account[half:], delta.ts[half:], delta.amount[half:], delta.tx_type[half:]))
    assert engine.history_info()["delta_events"] in (len(delta), 0)      # nothing was folded into the base (0: test knob)
    got = engine.get_velocity(q)
    bad = np.nonzero(got.view(np.uint8).reshape(len(q), -1) != want.view(np.uint8).reshape(len(q), -1))[0]
    assert bad.size == 0, [(names[int(q["account"][i])], int(q["ts"][i]) - T, int(q["seq_limit"][i])) for i in np.unique(bad)[:6]]
    engine.load_events(full, n_acc)
    assert engine.get_velocity(q).tobytes() == want.tobytes()


def test_delta_random_streams_and_compaction(engine, orc, pkg):
    """Time-ordered appends of uneven sizes onto a base, with forced (ts, amount) collisions and bursts: after every
    append the velocity blocks (incl. stream-mode queries cutting inside base and delta) equal the oracle on the
    concatenation.  With a small compaction threshold the delta is folded into the base on the way — same answers."""
    abi = pkg.abi
    n_acc, n = 120, 24000
    ev = synth.make_events(n, n_acc, seed=77, span=7 * 3600, collision_rate=0.08, zipf_s=0.9)
    order = np.argsort(ev.ts, kind="stable")
    ev = pkg.Events(ev.account[order], ev.ts[order], ev.amount[order], ev.tx_type[order])
    rng = np.random.default_rng(8)
    for threshold in (1 << 30, 3000):
        engine.set_compaction_threshold(threshold)
        try:
            cuts = [0, 9000, 9001, 9800, 13000, 13001, 20000, n]
            engine.load_events(cut(ev, 0, cuts[1], pkg), n_acc)
            if threshold == 3000:
                engine.prepare_ingest(8000)              # pre-allocated delta / spare arrays: same answers
            folded = False
            for lo, hi in zip(cuts[1:-1], cuts[2:]):
                engine.append_events(cut(ev, lo, hi, pkg))
                info = engine.history_info()
                assert info["n_events"] == hi
                folded |= info["delta_events"] < hi - cuts[1]
                q = np.zeros(4000, dtype=abi.query_dtype)
                pick = rng.integers(0, hi, q.shape[0])
                q["account"] = ev.account[pick]
                q["ts"] = ev.ts[pick] + rng.choice([0, 0, 1, 60, 61, 300, 1800, 1801, 3600, 3601, 90000], q.shape[0]).astype(np.uint32)
                q["seq_limit"] = abi.SEQ_ALL
                q["seq_limit"][::2] = rng.integers(0, hi + 1, q[::2].shape[0])
                got, want = engine.get_velocity(q), orc.velocity_batch(cut(ev, 0, hi, pkg), n_acc, q)
                assert got.tobytes() == want.tobytes(), f"threshold {threshold}, after appending [{lo}, {hi})"
            assert folded == (threshold == 3000) or os.environ.get("IGS_SORT_PAIRS")
        finally:
            engine.set_compaction_threshold(65536)


def test_append_overlaps_scoring_on_another_stream(engine, orc, pkg):
    """An append builds into arrays no kernel reads and swaps them in: scoring calls queued on another stream before,
    during and after appends return the history either without or with each batch — here every scoring launch
    is ordered against the appends by the host, so each must equal the oracle on exactly the prefix appended so far,
    while its kernels are still running when the next append starts."""
    import torch
    abi = pkg.abi
    lib = abi.load_library()
    import ctypes as C
    n_acc, n0, step, rounds = 3000, 200_000, 20_000, 6
    n = n0 + step * rounds
    ev = synth.make_events(n, n_acc, seed=91, span=6 * 3600, zipf_s=0.9)
    order = np.argsort(ev.ts, kind="stable")
    ev = pkg.Events(ev.account[order], ev.ts[order], ev.amount[order], ev.tx_type[order])
    players = synth.make_players(n_acc, seed=92, now=synth.T0 + 6 * 3600)
    model = synth.make_gbdt(n_trees=200, seed=93)
    engine.load_players(players)
    engine.load_gbdt(model)
    engine.load_events(cut(ev, 0, n0, pkg), n_acc)
    cfg = pkg.default_config(model_kind=abi.MODEL_GBDT)
    m = 320_000                                               # large enough to still be running when the append starts (and for k_score_ws)
    tx = synth.make_tx(m, n_acc, seed=94, t_lo=int(ev.ts[n0 - 1]), t_hi=int(ev.ts[-1]) + 100)
    dev = torch.device("cuda", 0)
    d_tx = torch.from_numpy(tx.view(np.uint8).reshape(m, -1)).to(dev)
    side = torch.cuda.Stream()
    outs = []
    for r in range(rounds):
        d_out = torch.zeros((m, 8), dtype=torch.uint8, device=dev)
        st = lib.igs_score_batch_device(engine.handle, C.byref(cfg), d_tx.data_ptr(), m, d_out.data_ptr(), None, None,
                                        side.cuda_stream)
        assert st == 0
        outs.append(d_out)
        engine.append_events(cut(ev, n0 + r * step, n0 + (r + 1) * step, pkg))     # while that kernel runs
    torch.cuda.synchronize()
    pick = np.random.default_rng(3).choice(m, 3000, replace=False)
    for r, d_out in enumerate(outs):
        got = np.frombuffer(d_out.cpu().numpy().tobytes(), dtype=abi.result_dtype)[pick]
        want = orc.score_batch(cfg, model, players, cut(ev, 0, n0 + r * step, pkg), np.ascontiguousarray(tx[pick]))
        assert got.tobytes() == want.tobytes(), f"scoring launch {r} saw a different history"


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


def test_eviction_is_invisible_to_present_and_future_queries(engine, orc, pkg, late_reorder):
    """igs_evict_events(now) = the TTLs of the Redis keys: what it drops cannot influence any query at
    t >= now.  History with dormant accounts (dead for more than 7 days), accounts that paused for more than a
    day and came back, and accounts that never pause; with identifier columns.  Velocity blocks and unique
    counts at t >= now are bit-identical before / after eviction and equal to the oracle on the FULL history;
    appends afterwards still behave like the full history."""
    abi = pkg.abi
    rng = np.random.default_rng(23)
    n_acc = 600
    day = 86400
    now = synth.T0 + 30 * day
    rows = []
    for a in range(n_acc):
        kind = a % 4
        if kind == 0:      # dormant: last event more than 7 days before now
            ts = now - rng.integers(8 * day, 25 * day) - np.sort(rng.integers(0, 2 * day, rng.integers(1, 30)))
        elif kind == 1:    # old burst, pause of more than a day, recent chain
            old = now - rng.integers(5 * day, 20 * day) - rng.integers(0, 3600, rng.integers(1, 40))
            new = now - rng.integers(0, 20 * 3600, rng.integers(1, 40))
            ts = np.concatenate([old, new])
        elif kind == 2:    # active every few hours for weeks: one long 24h-gap chain
            ts = now - np.cumsum(rng.integers(600, 20 * 3600, rng.integers(20, 120)))
        else:              # last seen 2-6 days ago: chain keys dead, last_tx alive
            ts = now - rng.integers(2 * day, 6 * day) - rng.integers(0, 7200, rng.integers(1, 10))
        rows.append((np.full(ts.size, a), ts))
    # a future-dated event more than a day after the account's last one (clock skew between the ingest clock and
    # the `now` handed to evict): the break it forms lies beyond `now` and must not drop what precedes it
    rows.append((np.full(3, n_acc), np.array([now - 4000, now - 100, now + 2 * day])))
    n_acc += 1
    acct = np.concatenate([r[0] for r in rows]).astype(np.uint32)
    ts = np.concatenate([r[1] for r in rows]).astype(np.uint32)
    perm = rng.permutation(acct.size)
    acct, ts = acct[perm], ts[perm]
    n = acct.size
    ev = pkg.Events(acct, ts, rng.integers(1, 50, n).astype(np.int64) * 100, rng.integers(0, 3, n).astype(np.uint8),
                    device=rng.integers(0, 6, n).astype(np.uint64), ip=rng.integers(0, 9, n).astype(np.uint64))
    engine.load_events(ev, n_acc)
    q = np.zeros(6000, dtype=abi.query_dtype)
    q["account"] = rng.integers(0, n_acc, q.shape[0])
    q["ts"] = now + rng.choice([0, 1, 60, 3600, day, 3 * day, 7 * day, 8 * day], q.shape[0]) + rng.integers(0, 100, q.shape[0])
    q["ts"][:50] = now
    q["account"][:4], q["ts"][:4] = n_acc - 1, [now, now + 50, now + day, now + 3 * day]
    q["seq_limit"] = abi.SEQ_ALL
    v0, u0 = engine.get_velocity(q), engine.get_uniques(q)
    assert v0.tobytes() == orc.velocity_batch(ev, n_acc, q).tobytes()
    dropped = engine.evict_events(now)
    assert 0.2 * n < dropped < 0.9 * n, dropped
    v1, u1 = engine.get_velocity(q), engine.get_uniques(q)
    same = v0.copy()
    same["n_prior"] = v1["n_prior"]                       # the number of prior events is the one thing that shrinks
    assert v1.tobytes() == same.tobytes()
    assert (v1["n_prior"] <= v0["n_prior"]).all() and (v1["n_prior"] < v0["n_prior"]).any()
    assert u1.tobytes() == u0.tobytes()
    assert v1["n_prior"][:2].tolist() == [2, 2] and int(v1["time_since_last_tx"][0]) == 100
    assert engine.evict_events(now) == 0                 # idempotent
    # appends after the eviction: as if nothing had been dropped
    m = 3000
    more = pkg.Events(rng.integers(0, n_acc, m).astype(np.uint32), (now + np.sort(rng.integers(0, 3 * day, m))).astype(np.uint32),
                      rng.integers(1, 50, m).astype(np.int64) * 100, rng.integers(0, 3, m).astype(np.uint8),
                      device=rng.integers(0, 6, m).astype(np.uint64), ip=rng.integers(0, 9, m).astype(np.uint64))
    engine.append_events(more)
    full = pkg.Events(np.concatenate([ev.account, more.account]), np.concatenate([ev.ts, more.ts]),
                      np.concatenate([ev.amount, more.amount]), np.concatenate([ev.tx_type, more.tx_type]),
                      device=np.concatenate([ev.device, more.device]), ip=np.concatenate([ev.ip, more.ip]))
    q2 = q.copy()
    q2["ts"] = now + 3 * day + rng.integers(0, 5 * day, q.shape[0])
    want = orc.velocity_batch(full, n_acc, q2)
    got = engine.get_velocity(q2)
    want["n_prior"] = got["n_prior"]
    assert got.tobytes() == want.tobytes()
    assert engine.get_uniques(q2).tobytes() == orc.uniques_batch(full, n_acc, q2).tobytes()
    # a later eviction drops the accounts that went quiet meanwhile
    assert engine.evict_events(now + 20 * day) > 0
