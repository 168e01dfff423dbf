"""GPU: what include/igs.h promises about threads — one engine per device handle, calls on different slots of one
engine and on different engines may come from different host threads at the same time (the reference serves one
goroutine per RPC on a shared ScoringEngine, services/risk/cmd/main.go:133-164)."""
import ctypes as C
import threading

import numpy as np
import pytest

from tests import synth

pytestmark = pytest.mark.gpu


def test_two_engines_two_threads_and_two_slots(orc, pkg):
    abi = pkg.abi
    lib = abi.load_library()
    n_acc, n_ev, n = 4000, 60_000, 150_000
    players = [synth.make_players(n_acc, seed=11 + k, now=synth.T0 + 6 * 3600) for k in range(2)]
    events = [synth.make_events(n_ev, n_acc, seed=21 + k, span=6 * 3600) for k in range(2)]
    model = synth.make_gbdt(n_trees=150, seed=31)
    cfg = pkg.default_config(model_kind=abi.MODEL_GBDT)
    txs = [synth.make_tx(n, n_acc, seed=41 + k, t_lo=synth.T0 + 3600, t_hi=synth.T0 + 6 * 3600) for k in range(4)]
    engines = [pkg.Engine(device=0) for _ in range(2)]
    try:
        for k, eng in enumerate(engines):
            eng.load_gbdt(model)
            eng.load_players(players[k])
            eng.load_events(events[k], n_acc)
        outs = [np.zeros(n, dtype=abi.result_dtype) for _ in range(4)]
        pinned = []

        def pin(a):
            p = lib.igs_alloc_pinned(a.nbytes)
            buf = np.frombuffer((C.c_uint8 * a.nbytes).from_address(p), dtype=a.dtype).reshape(a.shape)
            buf[...] = a
            pinned.append(p)
            return buf
        h_tx = [pin(t) for t in txs]
        h_out = [pin(o) for o in outs]
        errors = []

        def worker(job):
            # jobs 0, 1: engine 0, slots 1 and 2 (asynchronous entry points); jobs 2, 3: engine 1, slots 1 and 2
            eng, slot = engines[job // 2], 1 + job % 2
            try:
                for _ in range(6):
                    st = lib.igs_score_batch_async(eng.handle, slot, C.byref(cfg), h_tx[job].ctypes.data, n, h_out[job].ctypes.data)
                    if st == 0:
                        st = lib.igs_wait(eng.handle, slot)
                    if st != 0:
                        errors.append((job, lib.igs_last_error(eng.handle).decode()))
                        return
            except Exception as ex:        # noqa: BLE001
                errors.append((job, repr(ex)))

        threads = [threading.Thread(target=worker, args=(j,)) for j in range(4)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=300)
        assert not errors, errors
        pick = np.random.default_rng(5).choice(n, 2000, replace=False)
        for job in range(4):
            k = job // 2
            want = orc.score_batch(cfg, model, players[k], events[k], np.ascontiguousarray(txs[job][pick]))
            assert h_out[job][pick].tobytes() == want.tobytes(), f"job {job} (engine {k}, slot {1 + job % 2})"
    finally:
        for p in pinned:
            lib.igs_free_pinned(p)
        for eng in engines:
            eng.close()
