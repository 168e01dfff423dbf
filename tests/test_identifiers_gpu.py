"""GPU: identifiers on the device — exact unique devices / IPs per account from the
event index (igs_load_events_ids, igs_get_uniques), the blacklist sets (igs_blacklist_*) and the fused
path with R3 / R4 / R8 fed from them (igs_score_batch_ids) — against the oracle.  All integer work:
bit-exact."""
import json
import os

import numpy as np
import pytest

from tests import synth

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture
def clean_engine(engine, pkg):
    """The session engine, returned to a state without identifiers / blacklists afterwards."""
    yield engine
    abi = pkg.abi
    for kind in (abi.BL_DEVICE, abi.BL_FINGERPRINT, abi.BL_IP):
        engine.blacklist_load(kind, np.zeros(0, np.uint64))
    engine.load_events(pkg.Events(np.zeros(0, np.uint32), np.zeros(0, np.uint32), np.zeros(0, np.int64),
                                  np.zeros(0, np.uint8)), 1)


def queries(pkg, tx):
    return synth.queries_of(tx)


def test_kat_uniques_on_device(clean_engine, orc, pkg):
    with open(os.path.join(GOLDEN, "kat_uniques.json")) as f:
        kat = json.load(f)
    rows = kat["rows"]
    n = len(rows)
    ev = pkg.Events(np.zeros(n, np.uint32), np.array([kat["t0"] + r["dt"] for r in rows], np.uint32),
                    np.full(n, 100, np.int64), np.zeros(n, np.uint8),
                    device=np.array([pkg.hash_id(r["device"]) for r in rows], np.uint64),
                    ip=np.array([pkg.hash_id(r["ip"]) for r in rows], np.uint64))
    clean_engine.load_events(ev, 1)
    q = np.zeros(n, dtype=pkg.abi.query_dtype)
    q["ts"] = ev.ts
    q["seq_limit"] = np.arange(n)
    u = clean_engine.get_uniques(q)
    for i, r in enumerate(rows):
        assert (int(u[i]["devices_24h"]), int(u[i]["ips_24h"])) == (r["devices"], r["ips"]), f"dt={r['dt']}: {r['why']}"


@pytest.mark.parametrize("span_h,zipf", [(6, 1.05), (96, 0.6)])
def test_uniques_match_oracle(clean_engine, orc, pkg, span_h, zipf):
    """Random multi-account history (Zipf accounts, bursts, empty identifiers, devices shared across
    accounts; the 96 h span crosses the 24 h TTL many times)."""
    n_acc, n_ev, n_q = 2000, 60000, 30000
    ev = synth.make_event_ids(synth.make_events(n_ev, n_acc, seed=31, span=span_h * 3600, zipf_s=zipf), seed=32)
    clean_engine.load_events(ev, n_acc)
    tx = synth.make_tx(n_q, n_acc, seed=33, t_lo=synth.T0, t_hi=synth.T0 + (span_h + 30) * 3600, zipf_s=zipf)
    tx["seq_limit"][::3] = np.arange(n_q)[::3] % n_ev          # a third of the queries see a prefix of the arrivals
    q = queries(pkg, tx)
    got = clean_engine.get_uniques(q)
    want = orc.uniques_batch(ev, n_acc, q)
    assert got.tobytes() == want.tobytes()
    assert want["devices_24h"].max() > 3 and want["ips_24h"].max() > 5      # R3 / R4 thresholds are crossed
    assert (want["devices_24h"] == 0).any()
    # the velocity block of an identifier-carrying history is the one of the plain history
    v1 = clean_engine.get_velocity(q)
    clean_engine.load_events(pkg.Events(ev.account, ev.ts, ev.amount, ev.tx_type), n_acc)
    assert clean_engine.get_velocity(q).tobytes() == v1.tobytes()


def test_uniques_hash_collisions_and_hot_account(clean_engine, orc, pkg):
    """Identifiers that collide in the 30-bit sort key stay distinct; one account with a long history."""
    ids = synth.colliding_ids(4)
    n = 5000
    rng = np.random.default_rng(5)
    acct = np.where(rng.random(n) < 0.8, 0, rng.integers(1, 4, n)).astype(np.uint32)
    ts = (synth.T0 + np.sort(rng.integers(0, 30 * 3600, n))).astype(np.uint32)
    dev = ids[rng.integers(0, 4, n)]
    ip = np.where(rng.random(n) < 0.5, ids[rng.integers(0, 2, n)], rng.integers(1, 40, n).astype(np.uint64))
    ev = pkg.Events(acct, ts, np.full(n, 5, np.int64), np.zeros(n, np.uint8), device=dev, ip=ip.astype(np.uint64))
    clean_engine.load_events(ev, 4)
    q = np.zeros(n, dtype=pkg.abi.query_dtype)
    q["account"], q["ts"], q["seq_limit"] = acct, ts, np.arange(n)
    got = clean_engine.get_uniques(q)
    want = orc.uniques_batch(ev, 4, q)
    assert got.tobytes() == want.tobytes()
    assert want["devices_24h"].max() == 4


def test_uniques_saturate_at_u16(clean_engine, orc, pkg):
    n = 70000
    ev = pkg.Events(np.zeros(n, np.uint32), (synth.T0 + np.arange(n) // 2).astype(np.uint32), np.ones(n, np.int64),
                    np.zeros(n, np.uint8), device=np.arange(1, n + 1, dtype=np.uint64),
                    ip=(np.arange(n, dtype=np.uint64) % np.uint64(3)) + np.uint64(1))
    clean_engine.load_events(ev, 1)
    q = np.zeros(3, dtype=pkg.abi.query_dtype)
    q["ts"] = [synth.T0 + 100, synth.T0 + 32767, synth.T0 + n]
    q["seq_limit"] = pkg.abi.SEQ_ALL
    got = clean_engine.get_uniques(q)
    assert got["devices_24h"].tolist() == [202, 65535, 65535]
    assert got["ips_24h"].tolist() == [3, 3, 3]


def test_append_with_identifiers_equals_bulk_load(clean_engine, orc, pkg):
    n_acc, n_ev = 500, 20000
    ev = synth.make_event_ids(synth.make_events(n_ev, n_acc, seed=41, span=40 * 3600), seed=42)
    order = np.argsort(ev.ts, kind="stable")
    ev = pkg.Events(ev.account[order], ev.ts[order], ev.amount[order], ev.tx_type[order],
                    device=ev.device[order], ip=ev.ip[order])
    cut = [0, 9000, 9001, 15000, n_ev]

    def part(a, b, with_ids=True):
        return pkg.Events(ev.account[a:b], ev.ts[a:b], ev.amount[a:b], ev.tx_type[a:b],
                          device=ev.device[a:b] if with_ids else None, ip=ev.ip[a:b] if with_ids else None)
    clean_engine.load_events(part(cut[0], cut[1]), n_acc)
    for a, b in zip(cut[1:-1], cut[2:]):
        clean_engine.append_events(part(a, b))
    tx = synth.make_tx(8000, n_acc, seed=43, t_lo=synth.T0, t_hi=synth.T0 + 50 * 3600)
    q = queries(pkg, tx)
    got = clean_engine.get_uniques(q)
    assert got.tobytes() == orc.uniques_batch(ev, n_acc, q).tobytes()
    # appending events without identifiers to such a history = events with empty identifiers
    extra = part(100, 600, with_ids=False)
    extra.ts = (extra.ts.astype(np.int64) + 45 * 3600).astype(np.uint32)
    clean_engine.append_events(extra)
    z = np.zeros(500, np.uint64)
    full = pkg.Events(np.concatenate([ev.account, extra.account]), np.concatenate([ev.ts, extra.ts]),
                      np.concatenate([ev.amount, extra.amount]), np.concatenate([ev.tx_type, extra.tx_type]),
                      device=np.concatenate([ev.device, z]), ip=np.concatenate([ev.ip, z]))
    assert clean_engine.get_uniques(q).tobytes() == orc.uniques_batch(full, n_acc, q).tobytes()
    assert clean_engine.get_velocity(q).tobytes() == orc.velocity_batch(full, n_acc, q).tobytes()


def test_identifier_state_errors(clean_engine, pkg):
    abi = pkg.abi
    plain = synth.make_events(100, 10, seed=1)
    clean_engine.load_events(plain, 10)
    with pytest.raises(pkg.IgsError) as ei:
        clean_engine.get_uniques(np.zeros(1, dtype=abi.query_dtype))
    assert ei.value.status == abi.ERR_STATE
    with pytest.raises(pkg.IgsError) as ei:
        clean_engine.append_events(synth.make_event_ids(plain))
    assert ei.value.status == abi.ERR_STATE and "identifiers" in str(ei.value)
    with pytest.raises(pkg.IgsError) as ei:
        clean_engine.blacklist_add(7, np.array([1], np.uint64))
    assert ei.value.status == abi.ERR_INVALID and "unknown blacklist type" in str(ei.value)


def test_blacklist_sets_match_oracle(clean_engine, orc, pkg):
    """SADD / SISMEMBER semantics incl. growth of the table across adds and replacement."""
    abi = pkg.abi
    rng = np.random.default_rng(9)
    pool = rng.integers(1, 1 << 63, 60000, dtype=np.uint64)
    dev, fp, ip = pool[:20000], pool[20000:20500], pool[30000:30003]
    clean_engine.blacklist_load(abi.BL_DEVICE, dev[:300])
    for a in range(300, 20000, 4925):                              # forces several table rebuilds
        clean_engine.blacklist_add(abi.BL_DEVICE, dev[a:a + 4925])
    clean_engine.blacklist_add(abi.BL_DEVICE, dev[:100])           # re-adding members changes nothing
    clean_engine.blacklist_load(abi.BL_FINGERPRINT, np.concatenate([fp, fp[:50], np.zeros(3, np.uint64)]))
    clean_engine.blacklist_load(abi.BL_IP, pool[40000:40100])
    clean_engine.blacklist_load(abi.BL_IP, ip)                     # replace
    n = 50000
    ids = np.zeros(n, dtype=abi.tx_ids_dtype)
    for k, members in (("device", dev), ("fingerprint", fp), ("ip", ip)):
        pick = rng.random(n)
        ids[k] = np.where(pick < 0.02, members[rng.integers(0, len(members), n)],
                          np.where(pick < 0.3, np.uint64(0), pool[rng.integers(0, len(pool), n)]))
    bl, keep = orc.make_blacklist(dev, fp, ip)
    got = clean_engine.blacklist_check(ids)
    want = orc.blacklist_check(bl, ids[:4000])
    assert np.array_equal(got[:4000], want)
    # full batch against numpy set membership
    ref = (np.isin(ids["device"], dev) & (ids["device"] != 0)) | (np.isin(ids["fingerprint"], fp) & (ids["fingerprint"] != 0)) \
        | (np.isin(ids["ip"], ip) & (ids["ip"] != 0))
    assert np.array_equal(got.astype(bool), ref)
    assert 0.02 < ref.mean() < 0.5
    clean_engine.blacklist_load(abi.BL_DEVICE, np.zeros(0, np.uint64))
    ref2 = (np.isin(ids["fingerprint"], fp) & (ids["fingerprint"] != 0)) | (np.isin(ids["ip"], ip) & (ids["ip"] != 0))
    assert np.array_equal(clean_engine.blacklist_check(ids).astype(bool), ref2)


@pytest.mark.parametrize("kind", ["mock", "gbdt"])
def test_fused_path_with_identifiers_matches_oracle(clean_engine, orc, pkg, kind):
    """igs_score_batch_ids on a history that carries identifiers: R3 / R4 from the index, R8 from the
    blacklist sets OR the tx's own bit; results and assembled features bit-exact against the oracle."""
    abi = pkg.abi
    n_players, n_events, n = 3000, 50000, 20000
    now = synth.T0 + 30 * 3600
    players = synth.make_players(n_players, seed=51, now=now)
    ev = synth.make_event_ids(synth.make_events(n_events, n_players, seed=52, span=30 * 3600), seed=53)
    tx = synth.make_tx(n, n_players, seed=54, t_lo=synth.T0 + 3600, t_hi=now, invalid_rate=0.02)
    tx["unique_devices_24h"] = 9            # must be ignored: the history carries identifiers
    tx["unique_ips_24h"] = 9
    rng = np.random.default_rng(55)
    ids = np.zeros(n, dtype=abi.tx_ids_dtype)
    ids["device"] = np.where(rng.random(n) < 0.1, np.uint64(0), ev.device[rng.integers(0, n_events, n)])
    ids["fingerprint"] = rng.integers(1, 5000, n).astype(np.uint64)
    ids["ip"] = ev.ip[rng.integers(0, n_events, n)]
    bl_dev = np.unique(ev.device[ev.device != 0])[:40]
    bl_fp = np.arange(1, 30, dtype=np.uint64)
    bl_ip = np.unique(ev.ip[ev.ip != 0])[:25]
    clean_engine.load_players(players)
    clean_engine.load_events(ev, n_players)
    clean_engine.blacklist_load(abi.BL_DEVICE, bl_dev)
    clean_engine.blacklist_load(abi.BL_FINGERPRINT, bl_fp)
    clean_engine.blacklist_load(abi.BL_IP, bl_ip)
    model = None
    if kind == "gbdt":
        model = synth.make_gbdt(n_trees=60, seed=56)
        clean_engine.load_gbdt(model)
    cfg = pkg.default_config(model_kind=abi.MODEL_GBDT if kind == "gbdt" else abi.MODEL_MOCK)
    bl, keep = orc.make_blacklist(bl_dev, bl_fp, bl_ip)
    got, gf, _ = clean_engine.score_batch(cfg, tx, want_features=True, ids=ids)
    want, wf = orc.score_batch_ids(cfg, model, players, ev, tx, bl=bl, tx_ids=ids, want_features=True)
    assert np.array_equal(gf.view(np.uint32), wf.view(np.uint32))
    assert got.tobytes() == want.tobytes()
    fr = [(want["flags"] & b != 0).mean() for b in (abi.R_MULTIPLE_DEVICES, abi.R_IP_COUNTRY_MISMATCH, abi.R_KNOWN_FRAUDSTER)]
    assert all(0.001 < f < 0.9 for f in fr), fr
    # without the ids array only the tx's own blacklist bit counts; uniques still come from the index
    got2 = clean_engine.score_batch(cfg, tx)
    want2 = orc.score_batch_ids(cfg, model, players, ev, tx)
    assert got2.tobytes() == want2.tobytes()
    assert (got2["flags"] & abi.R_KNOWN_FRAUDSTER != 0).sum() < (got["flags"] & abi.R_KNOWN_FRAUDSTER != 0).sum()


def test_large_batch_is_visited_in_account_order_with_identical_results(clean_engine, orc, pkg):
    """Batches of >= 262 144 tx go through the device-side account sort + row permutation: results, assembled
    features and velocity blocks must land in the caller's row order, bit-identical to the oracle (and to the
    same rows scored in a small, unsorted batch)."""
    abi = pkg.abi
    n_players, n_events, n = 4000, 60000, 300_000
    now = synth.T0 + 12 * 3600
    players = synth.make_players(n_players, seed=61, now=now)
    ev = synth.make_event_ids(synth.make_events(n_events, n_players, seed=62, span=12 * 3600), seed=63)
    tx = synth.make_tx(n, n_players, seed=64, t_lo=synth.T0 + 3600, t_hi=now, invalid_rate=0.01)
    tx["account"][::4999] = n_players + 7                    # beyond the tables
    rng = np.random.default_rng(65)
    ids = np.zeros(n, dtype=abi.tx_ids_dtype)
    ids["device"] = ev.device[rng.integers(0, n_events, n)]
    ids["fingerprint"] = rng.integers(1, 3000, n).astype(np.uint64)
    ids["ip"] = ev.ip[rng.integers(0, n_events, n)]
    bl_fp = np.arange(1, 40, dtype=np.uint64)
    clean_engine.load_players(players)
    clean_engine.load_events(ev, n_players)
    clean_engine.blacklist_load(abi.BL_FINGERPRINT, bl_fp)
    model = synth.make_gbdt(n_trees=40, seed=66)
    clean_engine.load_gbdt(model)
    cfg = pkg.default_config(model_kind=abi.MODEL_GBDT)
    # device-buffer entry point: the whole batch in one launch (the host-buffer call would cut it into chunks
    # below the sorting threshold)
    import ctypes as C
    import torch
    lib = abi.load_library()
    dev = torch.device("cuda", 0)
    d_tx = torch.from_numpy(tx.view(np.uint8).reshape(n, -1)).to(dev)
    d_ids = torch.from_numpy(ids.view(np.uint8).reshape(n, -1)).to(dev)
    d_out = torch.zeros((n, 8), dtype=torch.uint8, device=dev)
    d_f = torch.zeros((n, abi.NUM_FEATURES), dtype=torch.float32, device=dev)
    d_v = torch.zeros((n, 128), dtype=torch.uint8, device=dev)
    st = lib.igs_score_batch_ids_device(clean_engine.handle, C.byref(cfg), d_tx.data_ptr(), d_ids.data_ptr(), n,
                                        d_out.data_ptr(), d_f.data_ptr(), d_v.data_ptr(),
                                        torch.cuda.current_stream().cuda_stream)
    assert st == 0, lib.igs_last_error(clean_engine.handle).decode()
    torch.cuda.synchronize()
    got = d_out.cpu().numpy().view(abi.result_dtype).reshape(n)
    gf = d_f.cpu().numpy()
    gv = d_v.cpu().numpy().view(abi.velocity_dtype).reshape(n)
    # oracle on a sample of rows (the oracle scores a row from the row alone, so a sub-batch is a fair check)
    pick = np.sort(rng.choice(n, 6000, replace=False))
    bl, keep = orc.make_blacklist(None, bl_fp, None)
    want, wf = orc.score_batch_ids(cfg, model, players, ev, tx[pick], bl=bl, tx_ids=ids[pick], want_features=True)
    assert got[pick].tobytes() == want.tobytes()
    assert np.array_equal(gf[pick].view(np.uint32), wf.view(np.uint32))
    # every row against the unsorted small-batch path of the library itself
    for a in range(0, n, 100_000):
        small, sf, sv = clean_engine.score_batch(cfg, tx[a:a + 100_000], want_features=True, want_velocity=True,
                                                 ids=ids[a:a + 100_000])
        assert got[a:a + 100_000].tobytes() == small.tobytes()
        assert gv[a:a + 100_000].tobytes() == sv.tobytes()
        assert np.array_equal(gf[a:a + 100_000].view(np.uint32), sf.view(np.uint32))
