"""GPU: linked accounts / CheckBonusAbuse — the three igs_links_* stages and the table
they build against the oracle, on one shard and on several shards simulated with several engines on one
GPU (the exchange between stages is then a device-to-device copy; tools/links_multi_gpu.py runs the real
NCCL all-to-all on 2+ GPUs).  Integer work: bit-exact."""
import numpy as np
import pytest

from tests import synth

pytestmark = pytest.mark.gpu


def make_history(pkg, n_events, n_accounts, seed, shared_rate=0.08, pools=400, kiosk=True):
    """Events whose devices are mostly private to an account, sometimes drawn from a small shared pool
    (rings of 2..20 accounts), plus one kiosk device used by > IGS_LINK_MAX_ACCOUNTS accounts."""
    ev = synth.make_events(n_events, n_accounts, seed=seed, span=20 * 3600, zipf_s=0.7)
    rng = np.random.default_rng(seed)
    acct = ev.account.astype(np.uint64)
    private = synth.mix64(acct * np.uint64(977) + rng.integers(0, 2, n_events).astype(np.uint64)) | np.uint64(1)
    shared = synth.mix64(rng.integers(0, pools, n_events).astype(np.uint64) + np.uint64(10_000_000)) | np.uint64(1)
    pick = rng.random(n_events)
    dev = np.where(pick < shared_rate, shared, private)
    dev = np.where(pick > 0.93, np.uint64(0), dev)                       # events without a device
    if kiosk:
        dev[rng.random(n_events) < 0.02] = np.uint64(0xC0FFEE)           # public device
    return pkg.Events(ev.account, ev.ts, ev.amount, ev.tx_type, device=dev.astype(np.uint64),
                      ip=np.zeros(n_events, np.uint64))


def test_single_shard_links_match_oracle(engine, orc, pkg):
    abi, links = pkg.abi, pkg.links
    n_acc, n_ev = 3000, 60000
    ev = make_history(pkg, n_ev, n_acc, seed=81)
    players = synth.make_players(n_acc, seed=82)
    engine.load_players(players)
    engine.load_events(ev, n_acc)
    stats = links.build_links(engine)
    assert stats["pairs_sent"] == int((ev.device != 0).sum()) and stats["links_received"] > 0
    q = np.arange(n_acc + 5, dtype=np.uint32)                            # + a few beyond the table
    got = engine.get_links(q)
    want = orc.links(ev.device, ev.account.astype(np.uint64), q.astype(np.uint64))
    assert got.tobytes() == want.tobytes()
    assert want["n_linked"].max() >= 3 and (want["n_linked"] == 0).any()
    # the kiosk device links nobody: accounts whose only shared device is the kiosk have no links
    kiosk_accts = np.unique(ev.account[ev.device == 0xC0FFEE])
    assert kiosk_accts.size > abi.LINK_MAX_ACCOUNTS
    # CheckBonusAbuse
    ab = engine.check_bonus_abuse(q)
    assert ab.tobytes() == orc.check_bonus_abuse(players, want, q).tobytes()
    assert ab["is_abuser"].any() and not ab["is_abuser"].all()
    assert set(np.unique(ab["signals"])) >= {0, abi.ABUSE_MULTI_ACCOUNT}
    # the reference-facing mirror: RiskChecker.CheckBonusAbuse by account uuid
    import uuid
    sc = pkg.scoring
    d = sc.PlayerDirectory()
    uu = [uuid.UUID(int=30_000 + i) for i in range(n_acc)]
    for u in uu:
        d.add(u)
    mirror = sc.ScoringEngine(engine, d)
    ring = int(np.nonzero(want["n_linked"] >= 3)[0][0])
    is_abuser, score, signals, linked = mirror.CheckBonusAbuse(uu[ring])
    assert (is_abuser, np.float32(score)) == (bool(ab["is_abuser"][ring]), ab["abuse_score"][ring])
    assert "MULTI_ACCOUNT" in signals and "ACCOUNT_RING" in signals
    assert linked == [str(uu[int(g)]) for g in want["linked"][ring]]
    assert mirror.CheckBonusAbuse(uuid.UUID(int=1)) == (False, 0.0, [], [])      # unknown account
    names = (__import__("ctypes").c_char_p * 3)()
    k = abi.load_library().igs_abuse_signals(7, names)
    assert [names[i].decode() for i in range(k)] == ["BONUS_ONLY_PLAYER", "MULTI_ACCOUNT", "ACCOUNT_RING"]


def test_device_at_the_public_threshold(engine, orc, pkg):
    """A device on exactly IGS_LINK_MAX_ACCOUNTS accounts links them all; one more account and it links nobody."""
    abi, links = pkg.abi, pkg.links
    K = abi.LINK_MAX_ACCOUNTS
    n_acc = 2 * K + 10
    acct = np.concatenate([np.arange(K), np.arange(K, 2 * K + 1), [0, 0, 5]]).astype(np.uint32)
    dev = np.concatenate([np.full(K, 111), np.full(K + 1, 222), [111, 333, 333]]).astype(np.uint64)
    n = acct.size
    ev = pkg.Events(acct, np.full(n, synth.T0, np.uint32), np.ones(n, np.int64), np.zeros(n, np.uint8), device=dev,
                    ip=np.zeros(n, np.uint64))
    engine.load_events(ev, n_acc)
    links.build_links(engine)
    got = engine.get_links(np.arange(n_acc, dtype=np.uint32))
    want = orc.links(dev, acct.astype(np.uint64), np.arange(n_acc, dtype=np.uint64))
    assert got.tobytes() == want.tobytes()
    assert got["n_linked"][1] == K - 1 and got["n_linked"][K + 3] == 0
    assert got["n_linked"][0] == K - 1 and got["n_linked"][5] == K - 1    # 0-5 share 333 too: already linked
    assert got["linked"][0].tolist() == [1, 2, 3]


@pytest.mark.parametrize("n_shards", [2, 5])
def test_simulated_shards_match_oracle(orc, pkg, built_lib, n_shards):
    """Several shards on one GPU: one engine per shard, accounts placed by igs_shard_of, the exchange done as
    device-to-device copies in the order an all-to-all delivers (rank-major)."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    abi, links = pkg.abi, pkg.links
    lib = abi.load_library()
    n_acc_total, n_ev = 4000, 80000
    ev = make_history(pkg, n_ev, n_acc_total, seed=91, shared_rate=0.1)
    # global placement: account g -> (shard, dense index on that shard)
    shard_of = np.array([lib.igs_shard_of(int(g), 0, n_shards) for g in range(n_acc_total)], dtype=np.uint32)
    dense = np.zeros(n_acc_total, dtype=np.uint32)
    sizes = []
    for s in range(n_shards):
        idx = np.nonzero(shard_of == s)[0]
        dense[idx] = np.arange(idx.size)
        sizes.append(idx.size)
    gid = (shard_of.astype(np.uint64) << np.uint64(32)) | dense.astype(np.uint64)
    engines = [pkg.Engine(0) for _ in range(n_shards)]
    try:
        for s, eng in enumerate(engines):
            m = shard_of[ev.account] == s
            eng.load_events(pkg.Events(dense[ev.account[m]], ev.ts[m], ev.amount[m], ev.tx_type[m],
                                       device=ev.device[m], ip=ev.ip[m]), sizes[s])

        def all_to_all(ptrs, counts):
            """ptrs[r], counts[r][s]: rank r's grouped send buffer -> per-rank receive tensors (rank-major)."""
            outs = []
            for dst in range(n_shards):
                parts = []
                for src in range(n_shards):
                    off = int(counts[src][:dst].sum()) * 16
                    nb = int(counts[src][dst]) * 16
                    if nb:
                        buf = torch.as_tensor(links._DeviceBytes(ptrs[src], int(counts[src].sum()) * 16), device="cuda:0")
                        parts.append(buf[off:off + nb].clone())
                outs.append(torch.cat(parts) if parts else torch.empty(0, dtype=torch.uint8, device="cuda:0"))
            torch.cuda.synchronize()
            return outs

        st1 = [eng.links_pairs(s, n_shards) for s, eng in enumerate(engines)]
        recv1 = all_to_all([p for p, _ in st1], [c for _, c in st1])
        st2 = [eng.links_group(t.data_ptr() if t.numel() else 0, t.numel() // 16, n_shards) for eng, t in zip(engines, recv1)]
        recv2 = all_to_all([p for p, _ in st2], [c for _, c in st2])
        for s, (eng, t) in enumerate(zip(engines, recv2)):
            eng.links_finish(t.data_ptr() if t.numel() else 0, t.numel() // 16, s)
        want_all = orc.links(ev.device, gid[ev.account], gid)
        crossing = 0
        for s, eng in enumerate(engines):
            idx = np.nonzero(shard_of == s)[0]
            got = eng.get_links(np.arange(idx.size, dtype=np.uint32))
            assert got.tobytes() == want_all[idx].tobytes(), f"shard {s}"
            crossing += int(((got["linked"][:, 0] >> np.uint64(32)) != s)[got["n_linked"] > 0].sum())
        assert crossing > 0, "no link crossed a shard boundary: the test would not exercise the exchange"
    finally:
        for eng in engines:
            eng.close()


def test_peer_window_single_shard_and_overflow(engine, orc, pkg):
    """The peer-memory path with one shard (its own window is the only destination), and a window that is too
    small: nothing is lost silently, the consuming stage names the capacity that was needed."""
    abi, links = pkg.abi, pkg.links
    n_acc, n_ev = 2000, 40000
    ev = make_history(pkg, n_ev, n_acc, seed=83)
    engine.load_events(ev, n_acc)
    links.setup_peer_windows(engine, 0, 1, cap_pairs=n_ev, cap_links=400_000)
    for _ in range(2):                                                   # a rerun reuses the window
        links.build_links_peer(engine)
        q = np.arange(n_acc, dtype=np.uint32)
        want = orc.links(ev.device, ev.account.astype(np.uint64), q.astype(np.uint64))
        assert engine.get_links(q).tobytes() == want.tobytes()
    n_pairs = int((ev.device != 0).sum())
    links.setup_peer_windows(engine, 0, 1, cap_pairs=n_pairs - 1, cap_links=400_000)
    with pytest.raises(pkg.IgsError) as ei:
        links.build_links_peer(engine)
    assert ei.value.status == abi.ERR_NOMEM and f"{n_pairs} pairs arrived" in str(ei.value)
    links.setup_peer_windows(engine, 0, 1, cap_pairs=n_pairs, cap_links=10)
    with pytest.raises(pkg.IgsError) as ei:
        links.build_links_peer(engine)
    assert ei.value.status == abi.ERR_NOMEM and "links arrived" in str(ei.value)
    engine.load_events(synth.make_events(10, 10, seed=1), 10)           # no identifiers
    with pytest.raises(pkg.IgsError) as ei:
        engine.links_pairs_peer(0, 1)
    assert ei.value.status == abi.ERR_STATE


@pytest.mark.parametrize("n_shards", [2, 4])
def test_peer_windows_between_simulated_shards(orc, pkg, built_lib, n_shards):
    """Several engines on one GPU write straight into each other's windows (attached by address: same
    process) — the kernels and the slot claiming are those of the multi-GPU job, only the wire is missing."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    abi, links = pkg.abi, pkg.links
    lib = abi.load_library()
    n_acc_total, n_ev = 3000, 60000
    ev = make_history(pkg, n_ev, n_acc_total, seed=95, shared_rate=0.1)
    shard_of = np.array([lib.igs_shard_of(int(g), 0, n_shards) for g in range(n_acc_total)], dtype=np.uint32)
    dense = np.zeros(n_acc_total, dtype=np.uint32)
    sizes = []
    for s in range(n_shards):
        idx = np.nonzero(shard_of == s)[0]
        dense[idx] = np.arange(idx.size)
        sizes.append(idx.size)
    gid = (shard_of.astype(np.uint64) << np.uint64(32)) | dense.astype(np.uint64)
    engines = [pkg.Engine(0) for _ in range(n_shards)]
    try:
        ptrs = []
        for s, eng in enumerate(engines):
            m = shard_of[ev.account] == s
            eng.load_events(pkg.Events(dense[ev.account[m]], ev.ts[m], ev.amount[m], ev.tx_type[m],
                                       device=ev.device[m], ip=ev.ip[m]), sizes[s])
            ptrs.append(eng.links_window_create(n_ev, 600_000)[1])
        for eng in engines:
            for s, p in enumerate(ptrs):
                eng.links_window_attach(s, p)
        for step in ("links_window_reset", "links_pairs_peer", "links_group_peer", "links_finish_peer"):
            for s, eng in enumerate(engines):                            # every shard finishes a step before the next
                if step == "links_window_reset":
                    eng.links_window_reset()
                elif step == "links_finish_peer":
                    eng.links_finish_peer(s)
                else:
                    getattr(eng, step)(s, n_shards)
        want_all = orc.links(ev.device, gid[ev.account], gid)
        for s, eng in enumerate(engines):
            idx = np.nonzero(shard_of == s)[0]
            assert eng.get_links(np.arange(idx.size, dtype=np.uint32)).tobytes() == want_all[idx].tobytes(), f"shard {s}"
    finally:
        for eng in engines:
            eng.close()


def test_links_need_identifiers(engine, pkg):
    engine.load_events(synth.make_events(100, 10, seed=1), 10)
    with pytest.raises(pkg.IgsError) as ei:
        pkg.links.build_links(engine)
    assert ei.value.status == pkg.abi.ERR_STATE
