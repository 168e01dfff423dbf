"""risk.v1.RiskService shim: descriptors match proto/risk/v1/risk.proto, the codec maps
requests/responses like the reference's handler would, and a real gRPC round trip works.  CPU tests use a
stand-in engine (a test double that returns canned result records); the GPU test goes through the library."""
import uuid

import numpy as np
import pytest

from tests import synth

# (field name -> number) spot checks taken from proto/risk/v1/risk.proto
REQ_FIELDS = {"account_id": 1, "player_id": 2, "amount": 3, "transaction_type": 4, "currency": 5, "game_id": 6,
              "round_id": 7, "ip_address": 8, "device_id": 9, "fingerprint": 10, "user_agent": 11, "session_id": 12,
              "metadata": 13}
RESP_FIELDS = {"score": 1, "action": 2, "reason_codes": 3, "rule_score": 4, "ml_score": 5, "response_time_ms": 6,
               "features": 7}


def test_descriptors_match_the_reference_proto(pkg):
    gs = pkg.grpc_service
    d = gs.ScoreTransactionRequest.DESCRIPTOR
    assert d.full_name == "risk.v1.ScoreTransactionRequest"
    assert {f.name: f.number for f in d.fields} == REQ_FIELDS
    assert d.fields_by_name["amount"].type == d.fields_by_name["amount"].TYPE_INT64
    r = gs.ScoreTransactionResponse.DESCRIPTOR
    assert {f.name: f.number for f in r.fields} == RESP_FIELDS
    assert r.fields_by_name["ml_score"].type == r.fields_by_name["ml_score"].TYPE_FLOAT        # float on the wire
    assert r.fields_by_name["action"].enum_type.values_by_name["ACTION_BLOCK"].number == 3
    fv = gs.FeatureVector.DESCRIPTOR
    assert len(fv.fields) == 26 and fv.fields_by_name["bonus_only_player"].number == 26
    assert fv.fields_by_name["tx_sum_1h"].number == 4 and fv.fields_by_name["time_since_last_tx_sec"].number == 16
    assert gs.ScoreBatchRequest.DESCRIPTOR.fields_by_name["transactions"].number == 1
    assert gs.ScoreBatchResponse.DESCRIPTOR.fields_by_name["results"].number == 1
    # blacklist + LTV messages (risk.proto:94-182)
    a = gs.AddToBlacklistRequest.DESCRIPTOR
    assert {f.name: f.number for f in a.fields} == {"type": 1, "value": 2, "reason": 3, "created_by": 4, "expires_at": 5}
    assert a.fields_by_name["expires_at"].message_type.full_name == "google.protobuf.Timestamp"
    c = gs.CheckBlacklistRequest.DESCRIPTOR
    assert {f.name: f.number for f in c.fields} == {"device_id": 1, "fingerprint": 2, "ip_address": 3, "email": 4}
    assert {f.name: f.number for f in gs.BlacklistMatch.DESCRIPTOR.fields} == {"type": 1, "value": 2, "reason": 3, "created_at": 4}
    b = gs.CheckBonusAbuseResponse.DESCRIPTOR
    assert {f.name: f.number for f in b.fields} == {"is_abuser": 1, "abuse_score": 2, "signals": 3, "linked_accounts": 4}
    assert {f.name: f.number for f in gs.CheckBonusAbuseRequest.DESCRIPTOR.fields} == {"account_id": 1, "bonus_id": 2}
    l = gs.PredictLTVResponse.DESCRIPTOR
    assert {f.name: f.number for f in l.fields} == {"account_id": 1, "predicted_ltv": 2, "segment": 3, "churn_risk": 4,
                                                    "predicted_active_days": 5, "confidence": 6, "next_best_action": 7,
                                                    "predicted_at": 8}
    assert l.fields_by_name["segment"].enum_type.values_by_name["SEGMENT_CHURNING"].number == 5
    assert gs.GetPlayerSegmentResponse.DESCRIPTOR.fields_by_name["recommended_actions"].is_repeated


def test_wire_bytes_are_plain_proto3(pkg):
    """Hand-encoded bytes a Go client would send parse into the expected message."""
    gs = pkg.grpc_service
    acc = b"00000000-0000-0000-0000-0000000003e9"
    raw = (b"\x0a" + bytes([len(acc)]) + acc          # 1: account_id
           + b"\x18\xa1\x8d\x06"                      # 3: amount = 100001 (varint)
           + b"\x22\x07deposit"                       # 4: transaction_type
           + b"\x42\x071.2.3.4"                       # 8: ip_address
           + b"\x6a\x06\x0a\x01k\x12\x01v")            # 13: metadata {"k": "v"}
    m = gs.ScoreTransactionRequest.FromString(raw)
    assert (m.account_id, m.amount, m.transaction_type, m.ip_address) == (acc.decode(), 100001, "deposit", "1.2.3.4")
    assert dict(m.metadata) == {"k": "v"}
    sr = gs.to_score_request(m)
    assert sr.account_id == uuid.UUID(int=1001) and sr.amount == 100001 and sr.tx_type == "deposit" and sr.ip == "1.2.3.4"
    resp = gs.ScoreTransactionResponse(score=42, action=2, reason_codes=["ML_HIGH_RISK"], ml_score=0.75)
    assert resp.SerializeToString().startswith(b"\x08\x2a\x10\x02\x1a\x0cML_HIGH_RISK")


class FakeDeviceEngine:
    """Test double for igs.Engine: returns canned igs_result / velocity records."""

    def __init__(self, abi):
        self.abi = abi
        self.calls = []

    def blacklist_add(self, kind, ids):
        self.sets = getattr(self, "sets", {0: set(), 1: set(), 2: set()})
        self.sets[kind] |= set(int(x) for x in ids)

    def blacklist_check(self, ids):
        sets = getattr(self, "sets", {0: set(), 1: set(), 2: set()})
        return np.array([int(r["device"]) in sets[0] or int(r["fingerprint"]) in sets[1] or int(r["ip"]) in sets[2]
                         for r in ids], dtype=np.uint8)

    def check_bonus_abuse(self, accounts):
        out = np.zeros(accounts.shape[0], dtype=self.abi.abuse_dtype)
        out["abuse_score"], out["n_linked"], out["is_abuser"] = 0.8, 2, 1
        out["signals"] = self.abi.ABUSE_BONUS_ONLY | self.abi.ABUSE_MULTI_ACCOUNT
        return out

    def get_links(self, accounts):
        out = np.zeros(accounts.shape[0], dtype=self.abi.links_row_dtype)
        out["n_linked"] = 2
        out["linked"][:, 0], out["linked"][:, 1] = 4, (3 << 32) | 9      # one on this shard, one on shard 3
        return out

    def ltv_batch(self, x, rows):
        out = np.zeros(rows.shape[0], dtype=self.abi.ltv_result_dtype)
        out["predicted_ltv"], out["churn_risk"], out["confidence"] = 1234.5, 0.25, 0.5
        out["days_survival"], out["segment"], out["next_best_action"] = 77, self.abi.SEG_HIGH, 0
        return out

    def score_batch(self, cfg, tx, want_features=False, want_velocity=False):
        n = tx.shape[0]
        self.calls.append(tx.copy())
        out = np.zeros(n, dtype=self.abi.result_dtype)
        out["score"] = np.minimum(tx["amount"] // 1000, 100)
        out["action"] = np.where(out["score"] >= cfg.block_threshold, 3, np.where(out["score"] >= cfg.review_threshold, 2, 1))
        out["flags"] = np.where(tx["bits"] & self.abi.BIT_TOR, self.abi.R_VPN_DETECTED | self.abi.R_ML_HIGH_RISK, 0)
        out["ml_score"] = 0.25
        vel = np.zeros(n, dtype=self.abi.velocity_dtype)
        vel["tx_count_1h"] = 4
        vel["tx_sum_1h"] = 1000
        vel["time_since_last_tx"] = 17
        return (out, None, vel) if (want_features or want_velocity) else out


def make_service(pkg, device_engine, players=None):
    sc, gs = pkg.scoring, pkg.grpc_service
    d = sc.PlayerDirectory()
    ids = [uuid.UUID(int=1000 + i) for i in range(8)]
    for u in ids:
        d.add(u)
    eng = sc.ScoringEngine(device_engine, d, config=pkg.default_config(),
                           ip_intel=lambda ip: (False, False, ip == "tor"))
    x, rows = synth.make_ltv_inputs(8, seed=4)
    return gs.RiskServicer(eng, players=players, ltv_inputs=(x, rows)), ids


def test_grpc_round_trip_with_stand_in_engine(pkg, built_lib):
    gs, abi = pkg.grpc_service, pkg.abi
    players = synth.make_players(8, seed=3)
    servicer, ids = make_service(pkg, FakeDeviceEngine(abi), players)
    server = gs.serve(servicer, "127.0.0.1:0")
    client = gs.RiskClient(f"127.0.0.1:{server.bound_port}")
    try:
        req = gs.ScoreBatchRequest(transactions=[
            gs.ScoreTransactionRequest(account_id=str(ids[0]), amount=85_000, transaction_type="deposit", ip_address="tor"),
            gs.ScoreTransactionRequest(account_id=str(ids[1]), amount=55_000, transaction_type="bet"),
            gs.ScoreTransactionRequest(account_id="not-a-uuid", amount=1_000, transaction_type="withdraw")])
        resp = client.ScoreBatch(req)
        assert [r.score for r in resp.results] == [85, 55, 1]
        assert [r.action for r in resp.results] == [3, 2, 1]
        assert list(resp.results[0].reason_codes) == ["VPN_DETECTED", "ML_HIGH_RISK"]     # emission order
        assert resp.results[0].rule_score == 15 and resp.results[1].rule_score == 0
        f0 = resp.results[0].features
        assert (f0.tx_count_1h, f0.tx_sum_1h, f0.tx_avg_1h, f0.time_since_last_tx_sec, f0.is_tor) == (4, 1000, 250.0, 17, True)
        assert f0.total_deposits == int(players["total_deposits"][0]) and f0.deposit_count == int(players["deposit_count"][0])
        assert resp.results[2].features.total_deposits == 0                                # unknown account
        tx = servicer.engine.engine.calls[-1]
        assert tx["account"].tolist() == [0, 1, 0xFFFFFFFE] and tx["tx_type"].tolist() == [abi.TX_DEPOSIT, abi.TX_BET, abi.TX_WITHDRAW]
        one = client.ScoreTransaction(req.transactions[1])
        assert (one.score, one.action) == (55, 2)
        th = client.GetThresholds(gs.GetThresholdsRequest())
        assert (th.block_threshold, th.review_threshold) == (80, 50)
        up = client.UpdateThresholds(gs.UpdateThresholdsRequest(block_threshold=50, review_threshold=10))
        assert up.success and up.block_threshold == 50
        assert client.ScoreTransaction(req.transactions[1]).action == 3
        import grpc
        # blacklist management on the (stand-in) device sets
        assert client.AddToBlacklist(gs.AddToBlacklistRequest(type="device", value="dev-9", reason="chargeback")).success
        assert not client.AddToBlacklist(gs.AddToBlacklistRequest(type="email", value="a@b.c")).success    # redis_store.go:261
        cb = client.CheckBlacklist(gs.CheckBlacklistRequest(device_id="dev-9", fingerprint="dev-9", ip_address="dev-9"))
        assert cb.is_blacklisted and [(m.type, m.value) for m in cb.matches] == [("device", "dev-9")]
        assert not client.CheckBlacklist(gs.CheckBlacklistRequest(device_id="dev-8")).is_blacklisted
        assert not client.CheckBlacklist(gs.CheckBlacklistRequest()).is_blacklisted
        # LTV
        lt = client.PredictLTV(gs.PredictLTVRequest(account_id=str(ids[2])))
        assert (lt.account_id, lt.segment, lt.predicted_active_days) == (str(ids[2]), 2, 77)
        assert lt.predicted_ltv == 1234.5 and lt.predicted_at.seconds > 0 and lt.next_best_action
        sg = client.GetPlayerSegment(gs.GetPlayerSegmentRequest(account_id=str(ids[2])))
        assert sg.segment == 2 and list(sg.recommended_actions) == [lt.next_best_action]
        import grpc
        with pytest.raises(grpc.RpcError) as ei:
            client.PredictLTV(gs.PredictLTVRequest(account_id=str(uuid.UUID(int=1))))
        assert ei.value.code() == grpc.StatusCode.NOT_FOUND
        ba = client.CheckBonusAbuse(gs.CheckBonusAbuseRequest(account_id=str(ids[1])))
        assert ba.is_abuser and abs(ba.abuse_score - 0.8) < 1e-6
        assert list(ba.signals) == ["BONUS_ONLY_PLAYER", "MULTI_ACCOUNT"] and list(ba.linked_accounts) == [str(ids[4]), "3:9"]
        gf = client.GetFeatures(gs.GetFeaturesRequest(account_id=str(ids[0])))
        assert gf.account_id == str(ids[0]) and gf.computed_at.seconds > 0
        assert (gf.features.tx_count_1h, gf.features.time_since_last_tx_sec) == (4, 17)
        assert gf.features.total_deposits == int(players["total_deposits"][0])
        raw = client.channel.unary_unary(f"/{gs.SERVICE}/NoSuchMethod", request_serializer=lambda b: b,
                                         response_deserializer=lambda b: b)
        with pytest.raises(grpc.RpcError) as ei:
            raw(b"")
        assert ei.value.code() == grpc.StatusCode.UNIMPLEMENTED
    finally:
        client.close()
        server.stop(0)


@pytest.mark.gpu
def test_grpc_score_batch_on_the_gpu(engine, orc, pkg):
    sc, gs, abi = pkg.scoring, pkg.grpc_service, pkg.abi
    n_players = 50
    now = int(__import__("time").time())
    players = synth.make_players(n_players, seed=61, now=now)
    ev = synth.make_events(3000, n_players, seed=62, span=3600, zipf_s=0.4)
    ev.ts = (ev.ts.astype(np.int64) - synth.T0 + now - 3600).astype(np.uint32)     # the last hour before now
    engine.load_players(players)
    engine.load_events(ev, n_players)
    d = sc.PlayerDirectory()
    ids = [uuid.UUID(int=5000 + i) for i in range(n_players)]
    for u in ids:
        d.add(u)
    eng = sc.ScoringEngine(engine, d, config=pkg.default_config(model_kind=abi.MODEL_MOCK))
    servicer = gs.RiskServicer(eng, players=players)
    server = gs.serve(servicer, "127.0.0.1:0")
    client = gs.RiskClient(f"127.0.0.1:{server.bound_port}")
    try:
        rng = np.random.default_rng(1)
        msgs = [gs.ScoreTransactionRequest(account_id=str(ids[int(rng.integers(0, n_players))]),
                                           amount=int(rng.integers(1, 300000)),
                                           transaction_type=["bet", "deposit", "withdraw"][int(rng.integers(0, 3))])
                for _ in range(500)]
        resp = client.ScoreBatch(gs.ScoreBatchRequest(transactions=msgs))
        assert len(resp.results) == 500
        # the same requests straight through the oracle at the timestamps the handler used
        # (wall clock moved at most a few seconds; compare fields that do not depend on "now")
        for m, r in zip(msgs, resp.results):
            acc = d.get(uuid.UUID(m.account_id))
            assert r.features.total_deposits == int(players["total_deposits"][acc])
            assert r.features.tx_count_1h >= r.features.tx_count_5m >= r.features.tx_count_1m
            assert 0 <= r.score <= 100 and r.action in (1, 2, 3)
            assert r.rule_score == sum(w for code, w in (("HIGH_VELOCITY", 20), ("NEW_ACCOUNT_LARGE_TX", 30),
                                                         ("RAPID_DEPOSIT_WITHDRAW", 25), ("BONUS_ABUSE", 20))
                                       if code in r.reason_codes) or r.rule_score == 100
        assert any(r.features.tx_count_1h > 0 for r in resp.results)
    finally:
        client.close()
        server.stop(0)


@pytest.mark.gpu
def test_grpc_blacklist_and_ltv_on_the_gpu(engine, orc, pkg):
    """AddToBlacklist -> ScoreTransaction fires KNOWN_FRAUDSTER through the device-side sets; CheckBlacklist
    names the matching list; PredictLTV / GetPlayerSegment return the LTV model's result for the account."""
    sc, gs, abi = pkg.scoring, pkg.grpc_service, pkg.abi
    n_players = 20
    now = int(__import__("time").time())
    players = synth.make_players(n_players, seed=71, now=now)
    engine.load_players(players)
    engine.load_events(pkg.Events(np.zeros(0, np.uint32), np.zeros(0, np.uint32), np.zeros(0, np.int64), np.zeros(0, np.uint8),
                                  device=np.zeros(0, np.uint64), ip=np.zeros(0, np.uint64)), n_players)
    mlp = synth.make_mlp(seed=72)
    engine.load_mlp(mlp)
    x, rows = synth.make_ltv_inputs(n_players, seed=73)
    d = sc.PlayerDirectory()
    ids = [uuid.UUID(int=8000 + i) for i in range(n_players)]
    for u in ids:
        d.add(u)
    eng = sc.ScoringEngine(engine, d, config=pkg.default_config(model_kind=abi.MODEL_NONE), device_identifiers=True)
    server = gs.serve(gs.RiskServicer(eng, players=players, ltv_inputs=(x, rows)), "127.0.0.1:0")
    client = gs.RiskClient(f"127.0.0.1:{server.bound_port}")
    try:
        req = gs.ScoreTransactionRequest(account_id=str(ids[3]), amount=500, transaction_type="bet", device_id="phone-1",
                                         fingerprint="fp-1", ip_address="198.51.100.9")
        assert "KNOWN_FRAUDSTER" not in client.ScoreTransaction(req).reason_codes
        assert client.AddToBlacklist(gs.AddToBlacklistRequest(type="fingerprint", value="fp-1")).success
        r = client.ScoreTransaction(req)
        assert "KNOWN_FRAUDSTER" in r.reason_codes and r.rule_score >= 50
        cb = client.CheckBlacklist(gs.CheckBlacklistRequest(device_id="phone-1", fingerprint="fp-1", ip_address="198.51.100.9"))
        assert cb.is_blacklisted and [(m.type, m.value) for m in cb.matches] == [("fingerprint", "fp-1")]
        assert not client.CheckBlacklist(gs.CheckBlacklistRequest(device_id="fp-1")).is_blacklisted   # wrong list
        # uniques echo: after two completed transactions with different devices the next score sees 2 devices
        for k in range(2):
            eng.UpdateFeatures([sc.TransactionEvent(ids[3], 100, "bet", ip="198.51.100.9", device_id=f"phone-{k}", timestamp=now - 5)])
        f = client.ScoreTransaction(req).features
        assert (f.unique_devices_24h, f.unique_ips_24h, f.tx_count_1m) == (2, 1, 1)   # equal (ts, amount): one ZSET member
        # LTV for one account == the batch call for that row
        want = engine.ltv_batch(x[5:6], rows[5:6])[0]
        lt = client.PredictLTV(gs.PredictLTVRequest(account_id=str(ids[5])))
        assert lt.predicted_ltv == np.float32(want["predicted_ltv"]) and lt.segment == int(want["segment"])
        assert lt.churn_risk == np.float32(want["churn_risk"]) and lt.predicted_active_days == int(want["days_survival"])
        sg = client.GetPlayerSegment(gs.GetPlayerSegmentRequest(account_id=str(ids[5])))
        assert sg.segment == lt.segment and list(sg.recommended_actions) == [lt.next_best_action]
    finally:
        client.close()
        server.stop(0)
        for kind in (abi.BL_DEVICE, abi.BL_FINGERPRINT, abi.BL_IP):
            engine.blacklist_load(kind, np.zeros(0, np.uint64))
        engine.load_events(pkg.Events(np.zeros(0, np.uint32), np.zeros(0, np.uint32), np.zeros(0, np.int64),
                                      np.zeros(0, np.uint8)), 1)


@pytest.mark.gpu
def test_grpc_concurrent_rpcs_do_not_share_staging(engine, orc, pkg):
    """The server runs a pool of worker threads and every synchronous engine call stages through slot 0
    (include/igs.h, Threading).  Eight client threads fire ScoreBatch requests of different sizes while another
    keeps adding to the device blacklist (its growth path re-allocates the set a scoring kernel probes): every
    response must be the one for ITS request — compared with the oracle per request."""
    import threading
    sc, gs, abi = pkg.scoring, pkg.grpc_service, pkg.abi
    n_players = 64
    players = synth.make_players(n_players, seed=81, now=synth.T0 + 3600)
    engine.load_players(players)
    engine.load_events(pkg.Events(np.zeros(0, np.uint32), np.zeros(0, np.uint32), np.zeros(0, np.int64), np.zeros(0, np.uint8)),
                       n_players)
    for kind in (abi.BL_DEVICE, abi.BL_FINGERPRINT, abi.BL_IP):
        engine.blacklist_load(kind, np.zeros(0, np.uint64))
    d = sc.PlayerDirectory()
    ids = [uuid.UUID(int=9000 + i) for i in range(n_players)]
    for u in ids:
        d.add(u)
    cfg = pkg.default_config(model_kind=abi.MODEL_MOCK)
    eng = sc.ScoringEngine(engine, d, config=cfg)
    server = gs.serve(gs.RiskServicer(eng, players=players), "127.0.0.1:0", max_workers=8)
    errors = []

    def scorer(k):
        client = gs.RiskClient(f"127.0.0.1:{server.bound_port}")
        rng = np.random.default_rng(100 + k)
        try:
            for rep in range(12):
                m = int(rng.integers(1, 400))
                acct = rng.integers(0, n_players, m)
                amount = rng.integers(1, 400000, m)
                typ = rng.integers(0, 3, m)
                msgs = [gs.ScoreTransactionRequest(account_id=str(ids[int(a)]), amount=int(v),
                                                   transaction_type=["bet", "deposit", "withdraw"][int(t)])
                        for a, v, t in zip(acct, amount, typ)]
                resp = client.ScoreBatch(gs.ScoreBatchRequest(transactions=msgs))
                # no history, no identifiers: the result depends on (player row, amount, type) only, except the
                # account age, which moves with the wall clock by whole days — recompute with the handler's day
                tx = np.zeros(m, dtype=abi.tx_dtype)
                tx["account"], tx["amount"] = acct, amount
                tx["tx_type"] = np.array([abi.TX_BET, abi.TX_DEPOSIT, abi.TX_WITHDRAW], np.uint8)[typ]
                tx["ts"], tx["seq_limit"], tx["bits"] = int(__import__("time").time()), abi.SEQ_ALL, abi.BIT_VALID
                want = orc.score_batch(cfg, None, players, pkg.Events(np.zeros(0, np.uint32), np.zeros(0, np.uint32),
                                                                      np.zeros(0, np.int64), np.zeros(0, np.uint8)), tx)
                got = [(r.score, r.action, r.rule_score) for r in resp.results]
                exp = [(int(w["score"]), int(w["action"]), 0) for w in want]
                if len(got) != m or any(g[:2] != e[:2] for g, e in zip(got, exp)):
                    errors.append((k, rep, m))
        except Exception as ex:            # noqa: BLE001
            errors.append((k, repr(ex)))
        finally:
            client.close()

    def blacklister():
        client = gs.RiskClient(f"127.0.0.1:{server.bound_port}")
        try:
            for i in range(300):
                assert client.AddToBlacklist(gs.AddToBlacklistRequest(type="device", value=f"bad-{i}")).success
        except Exception as ex:            # noqa: BLE001
            errors.append(("blacklist", repr(ex)))
        finally:
            client.close()

    threads = [threading.Thread(target=scorer, args=(k,)) for k in range(8)] + [threading.Thread(target=blacklister)]
    try:
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=120)
        assert not errors, errors[:5]
        hit = engine.blacklist_check(np.array([(pkg.hash_id("bad-299"), 0, 0)], dtype=abi.tx_ids_dtype))
        assert hit.tolist() == [1]
    finally:
        server.stop(0)
        engine.blacklist_load(abi.BL_DEVICE, np.zeros(0, np.uint64))
