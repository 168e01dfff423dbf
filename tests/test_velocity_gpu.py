"""GPU: event index (radix sort + segmented scans) and the velocity query kernel vs the oracle.
Everything here is integer work: counts, sums, gaps must be bit-exact."""
import json
import os

import numpy as np
import pytest

from tests import synth
from tests.test_oracle_kat import load

pytestmark = pytest.mark.gpu


def assert_velocity_equal(got, want):
    assert got.tobytes() == want.tobytes(), _first_diff(got, want)


def _first_diff(got, want):
    for k in ("n_prior", "tx_count_1m", "tx_count_5m", "tx_count_1h", "tx_sum_1h", "time_since_last_tx",
              "session_duration"):
        bad = np.nonzero(got[k] != want[k])[0]
        if bad.size:
            i = bad[0]
            return f"{k} differs at query {i}: got {got[k][i]} want {want[k][i]} ({bad.size} rows)"
    bad = np.nonzero((got["win"]["count"] != want["win"]["count"]).any(axis=(1, 2))
                     | (got["win"]["sum"] != want["win"]["sum"]).any(axis=(1, 2)))[0]
    return f"extension windows differ at {bad[:5]} ({bad.size} rows)"


def test_kat_set_b_on_device(engine, pkg):
    abi = pkg.abi
    kat = load("kat_velocity.json")
    rows = kat["rows"]
    n = len(rows)
    ev = pkg.Events(np.zeros(n, np.uint32), np.array([kat["t0"] + r["dt"] for r in rows], np.uint32),
                    np.array([r["amount"] for r in rows], np.int64), np.zeros(n, np.uint8))
    engine.load_events(ev, 1)
    q = np.zeros(n, dtype=abi.query_dtype)
    q["ts"] = ev.ts
    q["seq_limit"] = np.arange(n)
    v = engine.get_velocity(q)
    for i, r in enumerate(rows):
        got = tuple(int(v[i][k]) for k in ("tx_count_1m", "tx_count_5m", "tx_count_1h", "tx_sum_1h",
                                           "time_since_last_tx", "session_duration"))
        assert got == (r["c1m"], r["c5m"], r["c1h"], r["sum1h"], r["tslt"], r["sess"]), f"dt={r['dt']}"


def _random_queries(abi, rng, ev, n_accounts, nq):
    q = np.zeros(nq, dtype=abi.query_dtype)
    n = len(ev)
    pick = rng.integers(0, max(n, 1), size=nq)
    q["account"] = ev.account[pick] if n else 0
    q["ts"] = (ev.ts[pick] if n else synth.T0) + rng.choice([0, 0, 1, 59, 60, 61, 299, 300, 1800, 1801, 3599, 3600,
                                                             3601, 86400, 86401, 604800, 604801, 5_000_000],
                                                            size=nq).astype(np.uint32)
    q["seq_limit"] = abi.SEQ_ALL
    sl = rng.random(nq) < 0.3
    q["seq_limit"][sl] = rng.integers(0, max(n, 1) + 1, size=int(sl.sum()))
    # accounts with no events, and account ids beyond the index
    q["account"][rng.random(nq) < 0.05] = rng.integers(0, n_accounts)
    q["account"][rng.random(nq) < 0.01] = n_accounts + 5
    q["ts"][rng.random(nq) < 0.02] = 5      # long before any event
    return q


@pytest.mark.parametrize("n_events,n_accounts", [(0, 7), (1, 1), (2047, 300), (2048, 300), (2049, 300),
                                                 (4096, 3), (20000, 500), (60000, 20000)])
def test_velocity_matches_oracle(engine, orc, pkg, n_events, n_accounts):
    abi = pkg.abi
    rng = np.random.default_rng(n_events + 13)
    ev = synth.make_events(n_events, n_accounts, seed=n_events + 1, span=6 * 3600, collision_rate=0.05)
    engine.load_events(ev, n_accounts)
    q = _random_queries(abi, rng, ev, n_accounts, 4000)
    got = engine.get_velocity(q)
    want = orc.velocity_batch(ev, n_accounts, q)
    assert_velocity_equal(got, want)
    if n_events >= 20000:
        assert (want["tx_count_1m"] > 0).any() and (want["session_duration"] > 0).any()
        assert (want["win"]["count"][:, :, 0] > 0).any()


def test_dedupe_and_ties_stress(engine, orc, pkg):
    """One account, every event in the same second with few distinct amounts; queries cut the
    group at every arrival position (ZSET member dedupe)."""
    abi = pkg.abi
    n = 3000
    rng = np.random.default_rng(3)
    ts = np.full(n, synth.T0 + 100, np.uint32)
    ts[n // 2:] += 1
    ev = pkg.Events(np.zeros(n, np.uint32), ts, rng.integers(1, 6, size=n).astype(np.int64) * 100,
                    rng.integers(0, 3, size=n).astype(np.uint8))
    engine.load_events(ev, 1)
    q = np.zeros(n + 1, dtype=abi.query_dtype)
    q["ts"][:n] = ts
    q["ts"][n] = ts[-1]
    q["seq_limit"] = np.arange(n + 1)
    got = engine.get_velocity(q)
    want = orc.velocity_batch(ev, 1, q)
    assert_velocity_equal(got, want)
    assert int(want["tx_count_1m"].max()) <= 10      # at most 5 amounts x 2 seconds


def test_same_second_bursts_of_any_size(engine, orc, pkg, late_reorder):
    """Groups of events sharing (account, second) that are longer than the kernel's bounded backward walk
    (33 events) go through the per-group hash tables: sizes right at the boundary, thousands of events with
    few / many distinct amounts, the amount INT64_MIN (the tables' empty-slot marker), appended in pieces."""
    abi = pkg.abi
    rng = np.random.default_rng(17)
    I64_MIN = np.iinfo(np.int64).min
    parts = []          # (account, ts offset, amounts)
    parts.append((0, 100, rng.integers(1, 300, 6000)))                      # 6000 events, ~300 distinct amounts
    parts.append((0, 101, np.arange(5000, dtype=np.int64)))                 # all distinct
    parts.append((1, 100, np.full(4000, 77, np.int64)))                     # all equal
    parts.append((2, 50, np.concatenate([np.arange(33), [5]])))             # 34 events: dup is 29 back
    parts.append((3, 50, np.concatenate([[5], np.arange(100, 132), [5]])))  # 34 events: dup is 33 back (beyond the walk)
    parts.append((4, 50, np.arange(33)))                                    # 33 events: exactly the walk
    parts.append((5, 60, np.array([I64_MIN, 3, I64_MIN] + list(range(10, 50)) + [I64_MIN, 3], dtype=np.int64)))
    parts.append((6, 10, rng.integers(1, 50, 3000)))                        # plus ordinary traffic around it
    acct = np.concatenate([np.full(len(a), k, np.uint32) for k, _, a in parts] + [rng.integers(0, 8, 5000).astype(np.uint32)])
    ts = np.concatenate([np.full(len(a), synth.T0 + t, np.uint32) for _, t, a in parts]
                        + [(synth.T0 + rng.integers(0, 200, 5000)).astype(np.uint32)])
    amt = np.concatenate([np.asarray(a, np.int64) for _, _, a in parts] + [rng.integers(1, 20, 5000).astype(np.int64)])
    perm = rng.permutation(len(acct))                                       # arrival order is arbitrary
    ev = pkg.Events(acct[perm], ts[perm], amt[perm], np.zeros(len(acct), np.uint8))
    n_acc = 8
    engine.load_events(ev, n_acc)
    q = np.zeros(8 * 6, dtype=abi.query_dtype)
    q["account"] = np.repeat(np.arange(8), 6)
    q["ts"] = synth.T0 + np.tile([49, 50, 100, 101, 150, 3000], 8)
    q["seq_limit"] = abi.SEQ_ALL
    mid = np.zeros(64, dtype=abi.query_dtype)                               # queries that see part of a burst
    mid["account"] = rng.integers(0, 7, 64)
    mid["ts"] = synth.T0 + rng.choice([50, 60, 100, 101], 64)
    mid["seq_limit"] = rng.integers(0, len(acct), 64)
    q = np.concatenate([q, mid])
    want = orc.velocity_batch(ev, n_acc, q)
    assert_velocity_equal(engine.get_velocity(q), want)
    assert want["tx_count_1m"].max() >= 5000
    # the same history arriving in three appends
    cuts = [0, 7000, 7001, len(acct)]
    first = pkg.Events(ev.account[:cuts[1]], ev.ts[:cuts[1]], ev.amount[:cuts[1]], ev.tx_type[:cuts[1]])
    engine.load_events(first, n_acc)
    for a, b in zip(cuts[1:-1], cuts[2:]):
        engine.append_events(pkg.Events(ev.account[a:b], ev.ts[a:b], ev.amount[a:b], ev.tx_type[a:b]))
    assert_velocity_equal(engine.get_velocity(q), want)


def test_unsorted_arrival_and_wide_keys(engine, orc, pkg):
    """Events arrive in random time order over a 400-day span with sparse account ids (the sort
    must order by (account, ts) and keep arrival order among equal keys)."""
    abi = pkg.abi
    rng = np.random.default_rng(99)
    n, n_accounts = 30000, 3_000_000
    acct = rng.choice(rng.integers(0, n_accounts, size=400), size=n).astype(np.uint32)
    ts = (synth.T0 + rng.integers(0, 400 * 86400, size=n)).astype(np.uint32)
    dup = rng.random(n) < 0.2
    ts[dup] = ts[np.maximum(np.nonzero(dup)[0] - 1, 0)]
    acct[dup] = acct[np.maximum(np.nonzero(dup)[0] - 1, 0)]
    ev = pkg.Events(acct, ts, rng.integers(1, 4, size=n).astype(np.int64) * 50, rng.integers(0, 4, size=n).astype(np.uint8))
    engine.load_events(ev, n_accounts)
    q = _random_queries(abi, rng, ev, n_accounts, 5000)
    assert_velocity_equal(engine.get_velocity(q), orc.velocity_batch(ev, n_accounts, q))


def test_bad_account_rejected(engine, pkg):
    ev = pkg.Events(np.array([0, 9], np.uint32), np.array([1, 2], np.uint32), np.array([1, 1], np.int64),
                    np.zeros(2, np.uint8))
    with pytest.raises(pkg.IgsError) as ei:
        engine.load_events(ev, 5)
    assert ei.value.status == pkg.abi.ERR_INVALID


def test_c3_full_size_properties_and_sample(engine, orc, pkg):
    """BASELINE config 3: 10M events, 1M accounts, 1M queries in the last 24 h.  Size-independent
    properties on the full result + an oracle comparison on a sample of accounts."""
    abi = pkg.abi
    n_events, n_accounts, nq = 10_000_000, 1_000_000, 1_000_000
    ev = synth.make_events(n_events, n_accounts)
    engine.load_events(ev, n_accounts)
    tx = synth.make_tx(nq, n_accounts, zipf_s=1.05)
    q = synth.queries_of(tx)
    v = engine.get_velocity(q)
    # monotone windows
    assert (v["tx_count_1m"] <= v["tx_count_5m"]).all() and (v["tx_count_5m"] <= v["tx_count_1h"]).all()
    assert (v["tx_count_1h"] <= v["n_prior"]).all()
    c = v["win"]["count"]
    assert (c[:, :, 0] <= c[:, :, 1]).all() and (c[:, :, 1] <= c[:, :, 2]).all()
    assert ((c[:, 0, 2].astype(np.int64) + c[:, 1, 2]) <= v["n_prior"]).all()
    # checksum: querying every account after the last event must see every event exactly once
    allq = np.zeros(n_accounts, dtype=abi.query_dtype)
    allq["account"] = np.arange(n_accounts)
    allq["ts"] = int(ev.ts.max()) + 1
    allq["seq_limit"] = abi.SEQ_ALL
    va = engine.get_velocity(allq)
    assert int(va["n_prior"].astype(np.int64).sum()) == n_events
    assert np.array_equal(va["n_prior"], np.bincount(ev.account, minlength=n_accounts).astype(np.uint32))
    # 24h sums per type: sum over accounts == column sum of the events in the last 24 h
    last = ev.ts >= int(ev.ts.max()) + 1 - 86400
    for ty in (0, 1):
        m = last & (ev.tx_type == ty)
        assert int(va["win"]["sum"][:, ty, 2].sum()) == int(ev.amount[m].sum())
        assert int(va["win"]["count"][:, ty, 2].astype(np.int64).sum()) == int(m.sum())
    # oracle on a sample of accounts (cold and hot), restricted history
    rng = np.random.default_rng(1)
    sample_accounts = np.unique(np.concatenate([rng.integers(0, n_accounts, 300), np.arange(40)]))
    keep = np.isin(ev.account, sample_accounts)
    remap = np.full(n_accounts, 0, np.uint32)
    remap[sample_accounts] = np.arange(sample_accounts.size)
    sub = pkg.Events(remap[ev.account[keep]], ev.ts[keep], ev.amount[keep], ev.tx_type[keep])
    qs = q[np.isin(q["account"], sample_accounts)][:3000].copy()
    got = engine.get_velocity(qs)
    qs_sub = qs.copy()
    qs_sub["account"] = remap[qs["account"]]
    want = orc.velocity_batch(sub, sample_accounts.size, qs_sub)
    assert_velocity_equal(got, want)
