"""The reference-facing host mirror (igaming-platform_b200/scoring.py).  Marshalling is checked on
CPU; the end-to-end behaviour (Score / ScoreBatch / SetThresholds, reason strings in emission order)
on the GPU, reading like a test of scoring.ScoringEngine would."""
import uuid

import numpy as np
import pytest

from tests import synth


def test_marshal_requests_to_tx_records(pkg, built_lib):
    sc, abi = pkg.scoring, pkg.abi
    d = sc.PlayerDirectory()
    a, b = uuid.UUID(int=1), uuid.UUID(int=2)
    d.add(a), d.add(b)
    eng = sc.ScoringEngine(None, d, ip_intel=lambda ip: (ip == "10.0.0.1", False, ip == "tor"),
                           blacklist=lambda dev, fp, ip: dev == "bad-device",
                           uniques=lambda acc: (4, 6) if acc == b else (1, 1))
    tx = eng.marshal([sc.ScoreRequest(a, 1500, "deposit", ip="10.0.0.1", timestamp=1000),
                      sc.ScoreRequest(b, 99, "withdraw", ip="tor", device_id="bad-device", timestamp=2000),
                      sc.ScoreRequest(uuid.UUID(int=77), 5, "refund", timestamp=3000)])
    assert tx["account"].tolist() == [0, 1, 0xFFFFFFFE]
    assert tx["amount"].tolist() == [1500, 99, 5]
    assert tx["tx_type"].tolist() == [abi.TX_DEPOSIT, abi.TX_WITHDRAW, abi.TX_OTHER]
    assert tx["ts"].tolist() == [1000, 2000, 3000]
    assert tx["bits"].tolist() == [abi.BIT_VALID | abi.BIT_VPN, abi.BIT_VALID | abi.BIT_TOR | abi.BIT_BLACKLISTED,
                                   abi.BIT_VALID]
    assert tx["unique_devices_24h"].tolist() == [1, 4, 1] and tx["unique_ips_24h"].tolist() == [1, 6, 1]
    assert (tx["seq_limit"] == abi.SEQ_ALL).all()
    # nil ipIntel / blacklist are tolerated (engine.go:394, :469)
    plain = sc.ScoringEngine(None, d).marshal([sc.ScoreRequest(a, 1, "bet", ip="10.0.0.1", timestamp=5)])
    assert plain["bits"].tolist() == [abi.BIT_VALID]


def test_directory_shards_by_uuid(pkg, built_lib):
    d = pkg.scoring.PlayerDirectory(n_shards=8)
    shards = [d.shard_of(uuid.UUID(int=i * 0x9E3779B97F4A7C15 % (1 << 128))) for i in range(2000)]
    assert set(shards) == set(range(8))


@pytest.mark.gpu
def test_score_and_score_batch_like_the_reference(engine, orc, pkg):
    sc, abi = pkg.scoring, pkg.abi
    n_players = 64
    now = synth.T0 + 3600
    players = synth.make_players(n_players, seed=91, now=now)
    players["account_created_at"][0] = now - 2 * 86400          # new account
    players["account_created_at"][1] = now - 400 * 86400
    ev = synth.make_events(4000, n_players, seed=92, span=3600, zipf_s=0.5)
    engine.load_players(players)
    engine.load_events(ev, n_players)
    d = sc.PlayerDirectory()
    ids = [uuid.UUID(int=1000 + i) for i in range(n_players)]
    for u in ids:
        d.add(u)
    s = sc.ScoringEngine(engine, d, config=pkg.default_config(model_kind=abi.MODEL_MOCK),
                         ip_intel=lambda ip: (False, False, ip == "185.220.101.1"),
                         blacklist=lambda dev, fp, ip: fp == "fp-fraud",
                         uniques=lambda acc: (5, 7) if acc == ids[0] else (1, 1))
    reqs = [sc.ScoreRequest(ids[0], 250_000, "deposit", ip="185.220.101.1", fingerprint="fp-fraud", timestamp=now),
            sc.ScoreRequest(ids[1], 500, "bet", ip="8.8.8.8", timestamp=now),
            sc.ScoreRequest(uuid.UUID(int=5), 100, "bet", timestamp=now)]            # unknown account
    res = s.ScoreBatch(reqs)
    # the same batch through the oracle's per-request pipeline
    tx = s.marshal(reqs)
    want = orc.score_batch(s.config, None, players, ev, tx)
    for r, w in zip(res, want):
        assert (r.score, r.action) == (int(w["score"]), ["", "approve", "review", "block"][int(w["action"])])
        assert r.ml_score == float(w["ml_score"])
    r0 = res[0]
    assert r0.action == "block" and r0.score >= 80
    # emission order of applyRules: R2, R3, R4, R5, R8, then ML (engine.go:431-475, :285-287)
    assert [c for c in r0.reason_codes if c != "HIGH_VELOCITY" and c != "BONUS_ABUSE"] == [
        "NEW_ACCOUNT_LARGE_TX", "MULTIPLE_DEVICES", "IP_COUNTRY_MISMATCH", "VPN_DETECTED",
        "KNOWN_FRAUDSTER", "ML_HIGH_RISK"]
    assert r0.rule_score == 100
    single = s.Score(reqs[1])
    assert (single.score, single.action, single.reason_codes) == (res[1].score, res[1].action, res[1].reason_codes)
    assert single.features is not None and single.features.shape == (64,)
    # SetThresholds moves the action, not the score (engine.go:498-504)
    assert s.GetThresholds() == (80, 50)
    s.SetThresholds(101, 100)
    assert s.ScoreBatch(reqs)[0].action == "approve" and s.ScoreBatch(reqs)[0].score == r0.score
    s.SetThresholds(80, 50)
    text = s.ScoreWithExplanation(reqs[0])
    assert "Final Score:" in text and "KNOWN_FRAUDSTER (+50)" in text


@pytest.mark.gpu
def test_device_identifiers_stream_like_the_reference(engine, orc, pkg):
    """The reference's real-time loop with every identifier input on the device: Score a tx, then
    UpdateFeatures with it (main.go:219-224).  Devices / IPs accumulate per account until R3 / R4
    fire; AddToBlacklist makes R8 fire for a fingerprint.  Checked against the oracle's command-level
    Redis emulation of the same stream."""
    sc, abi = pkg.scoring, pkg.abi
    now = synth.T0 + 1000
    players = synth.make_players(4, seed=93, now=now)
    players["account_created_at"][:] = now - 400 * 86400
    engine.load_players(players)
    empty = pkg.Events(np.zeros(0, np.uint32), np.zeros(0, np.uint32), np.zeros(0, np.int64), np.zeros(0, np.uint8),
                       device=np.zeros(0, np.uint64), ip=np.zeros(0, np.uint64))
    engine.load_events(empty, 4)
    d = sc.PlayerDirectory()
    accts = [uuid.UUID(int=7000 + i) for i in range(4)]
    for u in accts:
        d.add(u)
    s = sc.ScoringEngine(engine, d, config=pkg.default_config(model_kind=abi.MODEL_NONE), device_identifiers=True)
    try:
        s.AddToBlacklist("fingerprint", "fp-bad")
        with pytest.raises(ValueError):
            s.AddToBlacklist("email", "x")
        L = orc.lib()
        redis = [L.orc_redis_new() for _ in accts]
        import ctypes as C
        fired = set()
        for step in range(9):
            t = now + 10 * step
            a = step % 2                                       # two accounts alternate
            dev, ip = f"dev-{a}-{step // 2}", f"10.0.{a}.{step}"
            fp = "fp-bad" if step == 8 else "fp-ok"
            r = s.Score(sc.ScoreRequest(accts[a], 500, "bet", ip=ip, device_id=dev, fingerprint=fp, timestamp=t))
            rt = orc.Realtime()
            L.orc_redis_get(redis[a], t, C.byref(rt))
            want = []
            if rt.unique_devices_24h > 3:
                want.append("MULTIPLE_DEVICES")
            if rt.unique_ips_24h > 5:
                want.append("IP_COUNTRY_MISMATCH")
            if fp == "fp-bad":
                want.append("KNOWN_FRAUDSTER")
            assert [c for c in r.reason_codes if c in ("MULTIPLE_DEVICES", "IP_COUNTRY_MISMATCH", "KNOWN_FRAUDSTER")] == want, step
            # features idx 5 / 6 = minmax(uniques) exactly as the emulation counts them
            assert r.features[5] == np.float32(min(rt.unique_devices_24h, 10)) / np.float32(10)
            fired |= set(want)
            s.UpdateFeatures([sc.TransactionEvent(accts[a], 500, "bet", ip=ip, device_id=dev, timestamp=t)])
            L.orc_redis_update_ids(redis[a], t, 500, orc.hash_id(dev), orc.hash_id(ip))
        assert fired == {"MULTIPLE_DEVICES", "KNOWN_FRAUDSTER"}
        for h in redis:
            L.orc_redis_free(h)
    finally:
        for kind in (abi.BL_DEVICE, abi.BL_FINGERPRINT, abi.BL_IP):
            engine.blacklist_load(kind, np.zeros(0, np.uint64))
        engine.load_events(pkg.Events(np.zeros(0, np.uint32), np.zeros(0, np.uint32), np.zeros(0, np.int64),
                                      np.zeros(0, np.uint8)), 1)
