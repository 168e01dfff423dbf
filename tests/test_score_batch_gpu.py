"""GPU: the fused path (compact tx -> velocity lookup -> feature assembly -> rules -> model ->
blend in one kernel) through igs_score_batch vs the oracle's per-request pipeline.
Flags / velocity / assembled features / ml / score / action all bit-exact."""
import numpy as np
import pytest

from tests import synth

pytestmark = pytest.mark.gpu


def check(engine, orc, pkg, cfg, model, players, ev, tx):
    got, gf, gv = engine.score_batch(cfg, tx, want_features=True, want_velocity=True)
    want, wf, wv = orc.score_batch(cfg, model, players, ev, tx, want_features=True, want_velocity=True)
    assert gv.tobytes() == wv.tobytes(), "velocity blocks differ"
    assert np.array_equal(gf.view(np.uint32), wf.view(np.uint32)), \
        f"assembled features differ in columns {np.unique(np.nonzero(gf.view(np.uint32) != wf.view(np.uint32))[1])}"
    assert np.array_equal(got["flags"], want["flags"]), "flags differ"
    assert np.max(np.abs(got["ml_score"].astype(np.float64) - want["ml_score"]), initial=0) <= 1e-5
    assert got.tobytes() == want.tobytes(), "results differ"
    return want


@pytest.mark.parametrize("kind", ["none", "mock", "gbdt"])
def test_fused_path_matches_oracle(engine, orc, pkg, kind):
    abi = pkg.abi
    n_players, n_events, n = 3000, 40000, 20000
    now = synth.T0 + 6 * 3600
    players = synth.make_players(n_players, seed=5, now=now)
    ev = synth.make_events(n_events, n_players, seed=6, span=6 * 3600, collision_rate=0.03)
    tx = synth.make_tx(n, n_players, seed=7, t_lo=synth.T0 + 3600, t_hi=now, invalid_rate=0.03)
    # a few rows from accounts beyond the player table / the index (feature store "fails" -> zeros)
    tx["account"][::997] = n_players + 3
    engine.load_players(players)
    engine.load_events(ev, n_players)
    model = None
    mk = {"none": abi.MODEL_NONE, "mock": abi.MODEL_MOCK, "gbdt": abi.MODEL_GBDT}[kind]
    if kind == "gbdt":
        model = synth.make_gbdt(n_trees=120, seed=8)
        # make the extension block and raw-magnitude features matter to the trees
        engine.load_gbdt(model)
    cfg = pkg.default_config(model_kind=mk)
    want = check(engine, orc, pkg, cfg, model, players, ev, tx)
    seen = np.bitwise_or.reduce(want["flags"])
    for bit in (abi.R_HIGH_VELOCITY, abi.R_NEW_ACCOUNT_LARGE_TX, abi.R_MULTIPLE_DEVICES,
                abi.R_IP_COUNTRY_MISMATCH, abi.R_VPN_DETECTED, abi.R_KNOWN_FRAUDSTER, abi.R_BONUS_ABUSE):
        assert seen & bit, f"rule bit {bit} never fired"


def test_fused_path_stream_mode(engine, orc, pkg):
    """Score-then-apply stream (the reference's real-time order): tx i sees the events that
    arrived before it (seq_limit = i) — the events ARE the transactions."""
    abi = pkg.abi
    n_players, n = 200, 6000
    players = synth.make_players(n_players, seed=15, now=synth.T0 + 3 * 3600)
    ev = synth.make_events(n, n_players, seed=16, span=3 * 3600, collision_rate=0.05, zipf_s=0.8)
    order = np.argsort(ev.ts, kind="stable")
    ev = pkg.Events(ev.account[order], ev.ts[order], ev.amount[order], ev.tx_type[order])
    tx = np.zeros(n, dtype=abi.tx_dtype)
    tx["account"], tx["ts"], tx["amount"], tx["tx_type"] = ev.account, ev.ts, ev.amount, ev.tx_type
    tx["seq_limit"] = np.arange(n)
    tx["unique_devices_24h"] = 1
    tx["unique_ips_24h"] = 1
    tx["bits"] = abi.BIT_VALID
    engine.load_players(players)
    engine.load_events(ev, n_players)
    cfg = pkg.default_config(model_kind=abi.MODEL_MOCK)
    want = check(engine, orc, pkg, cfg, None, players, ev, tx)
    assert (want["flags"] & abi.R_HIGH_VELOCITY).any(), "bursts should trip HIGH_VELOCITY"


def test_fused_path_without_state(pkg, orc):
    """No players, no events loaded: the reference degrades to zero features (engine.go:267-270)."""
    abi = pkg.abi
    with pkg.Engine(device=0) as fresh:
        tx = synth.make_tx(3000, 50, seed=3)
        cfg = pkg.default_config(model_kind=abi.MODEL_MOCK)
        got = fresh.score_batch(cfg, tx)
        empty = pkg.Events(np.zeros(0, np.uint32), np.zeros(0, np.uint32), np.zeros(0, np.int64), np.zeros(0, np.uint8))
        want = orc.score_batch(cfg, None, np.zeros(0, abi.player_row_dtype), empty, tx)
        assert got.tobytes() == want.tobytes()


def test_fused_equals_two_step_path(engine, pkg):
    """Path B == velocity query + host-visible features fed back through path A."""
    abi = pkg.abi
    n_players = 1000
    players = synth.make_players(n_players, seed=25)
    ev = synth.make_events(30000, n_players, seed=26, span=12 * 3600)
    tx = synth.make_tx(8000, n_players, seed=27, t_lo=synth.T0 + 6 * 3600, t_hi=synth.T0 + 12 * 3600)
    model = synth.make_gbdt(n_trees=75, seed=28)
    engine.load_players(players)
    engine.load_events(ev, n_players)
    engine.load_gbdt(model)
    cfg = pkg.default_config(model_kind=abi.MODEL_GBDT)
    got, feats, vel = engine.score_batch(cfg, tx, want_features=True, want_velocity=True)
    v2 = engine.get_velocity(synth.queries_of(tx))
    assert vel.tobytes() == v2.tobytes()
    rows = np.zeros(tx.shape[0], dtype=abi.rule_row_dtype)
    p = players[tx["account"]]
    rows["amount"] = tx["amount"]
    rows["total_deposits"] = p["total_deposits"]
    rows["total_withdrawals"] = p["total_withdrawals"]
    rows["tx_count_1m"] = vel["tx_count_1m"]
    rows["account_age_days"] = (tx["ts"].astype(np.int64) - p["account_created_at"]) // 86400
    rows["time_since_last_tx"] = vel["time_since_last_tx"]
    rows["deposit_count"] = p["deposit_count"]
    rows["unique_devices_24h"] = tx["unique_devices_24h"]
    rows["unique_ips_24h"] = tx["unique_ips_24h"]
    rows["tx_type"] = tx["tx_type"]
    bonus_only = (p["bonus_claim_count"] > 3) & (p["total_deposits"] < 5000)
    rows["bits"] = tx["bits"] | (bonus_only * abi.BIT_BONUS_ONLY).astype(np.uint8)
    again = engine.score_features(cfg, feats, rows)
    assert got.tobytes() == again.tobytes()


def test_c1_reference_config_mock_model(engine, orc, pkg):
    """BASELINE.json configs[0] — what the reference's own CPU run would score: 100 000 bet / deposit / withdraw
    transactions over 10 000 Zipf(1.1) accounts, rules + the only model the reference can run (mockPredict,
    ml/onnx_model.go:258-308), DefaultConfig.  Through igs_score_batch against the per-request restatement:
    velocity blocks, assembled features, flags, ml, score, action bit-exact on all 100 000 rows."""
    abi = pkg.abi
    n, n_acc, n_ev = 100_000, 10_000, 200_000
    now = synth.T0 + 48 * 3600
    players = synth.make_players(n_acc, seed=synth.MASTER_SEED + 1, now=now)
    ev = synth.make_events(n_ev, n_acc, seed=synth.MASTER_SEED + 2, zipf_s=1.1, collision_rate=0.01)
    tx = synth.make_tx(n, n_acc, seed=synth.MASTER_SEED + 3, zipf_s=1.1, t_lo=synth.T0 + 24 * 3600, t_hi=now)
    engine.load_players(players)
    engine.load_events(ev, n_acc)
    cfg = pkg.default_config()                       # DefaultConfig(): model_kind = IGS_MODEL_MOCK
    assert cfg.model_kind == abi.MODEL_MOCK
    want = check(engine, orc, pkg, cfg, None, players, ev, tx)
    types = np.bincount(tx["tx_type"], minlength=3) / n
    assert abs(types[abi.TX_BET] - 0.7) < 0.01 and abs(types[abi.TX_DEPOSIT] - 0.2) < 0.01
    assert len(np.unique(want["action"])) >= 2 and (want["flags"] & abi.R_HIGH_VELOCITY).any()


@pytest.mark.parametrize("kind", ["none", "mock", "gbdt"])
def test_large_batches_take_the_warp_specialised_kernel(engine, orc, pkg, kind):
    """Batches of at least 4 x 148 x 512 tx that want only the results run through k_score_ws (producer warps look up and
    assemble, walker warps walk the trees; the model image with compacted feature rows).  Ragged size, rows from
    unknown accounts, invalid rows, a model that splits on every feature index incl. the constant-zero ones, all three
    model kinds — bit-identical to the per-request restatement, and to k_score (which serves the same call when the
    assembled features are requested too)."""
    abi = pkg.abi
    n_players, n_events, n = 3000, 60000, 4 * 148 * 512 + 4321
    now = synth.T0 + 6 * 3600
    players = synth.make_players(n_players, seed=105, now=now)
    ev = synth.make_events(n_events, n_players, seed=106, span=6 * 3600, collision_rate=0.03)
    tx = synth.make_tx(n, n_players, seed=107, t_lo=synth.T0 + 3600, t_hi=now, invalid_rate=0.02)
    tx["account"][::1013] = n_players + 5
    engine.load_players(players)
    engine.load_events(ev, n_players)
    model = None
    if kind == "gbdt":
        model = synth.make_gbdt(n_trees=57, seed=108)          # 5 full chunks of 10 trees + a ragged one
        model.feat[:, ::3] = np.array([7, 8, 22, 42, 63, 50, 41, 0, 6, 9, 21, 23], dtype=np.uint8)[
            np.arange(model.feat[:, ::3].size).reshape(model.feat[:, ::3].shape) % 12]
        engine.load_gbdt(model)
    cfg = pkg.default_config(model_kind={"none": abi.MODEL_NONE, "mock": abi.MODEL_MOCK, "gbdt": abi.MODEL_GBDT}[kind])
    got = engine.score_batch(cfg, tx)                                          # k_score_ws
    via_k_score, _, _ = engine.score_batch(cfg, tx, want_features=True)        # k_score
    assert got.tobytes() == via_k_score.tobytes()
    pick = np.concatenate([np.arange(0, 3000), np.arange(n - 3000, n)])
    want = orc.score_batch(cfg, model, players, ev, np.ascontiguousarray(tx[pick]))
    assert got[pick].tobytes() == want.tobytes()
    assert len(np.unique(got["score"])) >= 3                  # (model nil: 0.4 x rule score never reaches review)


def test_large_pinned_batches_stream_in_behind_the_kernel(engine, orc, pkg):
    """igs_score_batch on pinned host buffers of >= 4 x 148 x 512 tx: one k_score_ws launch over the whole batch, the
    records copied in behind its start in 32 chunks with an arrival counter the producer warps wait on, results written
    straight into the pinned result buffer.  Ragged size; all rows bit-identical to the staged path (pageable buffers)
    and a sample to the oracle; then two such calls at once on two slots of the engine."""
    import ctypes as C
    import threading
    abi = pkg.abi
    lib = abi.load_library()
    n_players, n_events, n = 5000, 80000, 4 * 148 * 512 + 77_777
    now = synth.T0 + 6 * 3600
    players = synth.make_players(n_players, seed=205, now=now)
    ev = synth.make_events(n_events, n_players, seed=206, span=6 * 3600, collision_rate=0.03)
    engine.load_players(players)
    engine.load_events(ev, n_players)
    model = synth.make_gbdt(n_trees=40, seed=208)
    engine.load_gbdt(model)
    cfg = pkg.default_config(model_kind=abi.MODEL_GBDT)
    txs = [synth.make_tx(n, n_players, seed=207 + k, t_lo=synth.T0 + 3600, t_hi=now, invalid_rate=0.02) for k in range(2)]
    staged = [engine.score_batch(cfg, t) for t in txs]
    pinned = []

    def pin(a):
        p = lib.igs_alloc_pinned(a.nbytes)
        buf = np.frombuffer((C.c_uint8 * a.nbytes).from_address(p), dtype=a.dtype).reshape(a.shape)
        buf[...] = a
        pinned.append(p)
        return buf
    try:
        h_tx = [pin(t) for t in txs]
        h_out = [pin(np.full(n, 0xEE, dtype=np.uint8).repeat(8).view(abi.result_dtype)) for _ in txs]
        for rep in range(3):                 # the staging buffer is reused: stale records of the previous call must not be read
            k = rep % 2
            h_out[k].view(np.uint8)[...] = 0xEE
            st = lib.igs_score_batch(engine.handle, C.byref(cfg), h_tx[k].ctypes.data, n, h_out[k].ctypes.data, None, None)
            assert st == 0, lib.igs_last_error(engine.handle).decode()
            assert h_out[k].tobytes() == staged[k].tobytes()
        pick = np.concatenate([np.arange(0, 2000), np.arange(n - 2000, n)])
        want = orc.score_batch(cfg, model, players, ev, np.ascontiguousarray(txs[0][pick]))
        assert staged[0][pick].tobytes() == want.tobytes()
        errors = []

        def worker(k):
            for _ in range(4):
                h_out[k].view(np.uint8)[...] = 0xEE
                st = lib.igs_score_batch_async(engine.handle, 1 + k, C.byref(cfg), h_tx[k].ctypes.data, n, h_out[k].ctypes.data)
                if st == 0:
                    st = lib.igs_wait(engine.handle, 1 + k)
                if st != 0 or h_out[k].tobytes() != staged[k].tobytes():
                    errors.append((k, st, lib.igs_last_error(engine.handle).decode()))
                    return
        threads = [threading.Thread(target=worker, args=(k,)) for k in range(2)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=300)
        assert not errors, errors
    finally:
        for p in pinned:
            lib.igs_free_pinned(p)
