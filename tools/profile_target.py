#!/usr/bin/env python
"""Short single-scenario runs for ncu (a handful of launches each, no timing claims).

    python tools/profile_target.py fused      # 1M tx through igs_score_batch_device (C2 model, 500k players, 5M events)
    python tools/profile_target.py path_a     # 1M pre-assembled rows through igs_score_features_device
    python tools/profile_target.py ltv        # one 262 144-row chunk through igs_ltv_batch_device
    python tools/profile_target.py c3         # igs_load_events of 10M events / 1M accounts + 1M velocity queries
    python tools/profile_target.py append     # 100k-event appends onto the resident C3 history

Prints the CUDA-event time of the scenario's last repetition so a plain run documents itself.
"""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

ge.build_library()
pkg = ge.load_package()
abi = pkg.abi
lib = abi.load_library()
from tests import synth  # noqa: E402
import torch  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "fused"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream().cuda_stream
eng = pkg.Engine(0)


def timed(fn):
    ms = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    return ms


out = {"scenario": what}
if what in ("fused", "path_a"):
    n = 1_000_000
    cfg = pkg.default_config(model_kind=abi.MODEL_GBDT)
    eng.load_gbdt(synth.make_gbdt(500))
    d_out = torch.zeros((n, 8), dtype=torch.uint8, device=dev)
    if what == "fused":
        n_pl, n_ev = 500_000, 5_000_000
        eng.load_players(synth.make_players(n_pl))
        eng.load_events(synth.make_events(n_ev, n_pl), n_pl)
        tx = synth.make_tx(n, n_pl)
        d_tx = torch.from_numpy(tx.view(np.uint8).reshape(n, -1)).to(dev)
        out["ms"] = timed(lambda: lib.igs_score_batch_device(eng.handle, C.byref(cfg), d_tx.data_ptr(), n,
                                                             d_out.data_ptr(), None, None, stream))
    else:
        feats, rows = synth.make_feature_batch(n)
        d_f = torch.from_numpy(feats).to(dev)
        d_r = torch.from_numpy(rows.view(np.uint8).reshape(n, -1)).to(dev)
        out["ms"] = timed(lambda: eng.score_features_device(cfg, d_f.data_ptr(), d_r.data_ptr(), n, d_out.data_ptr(), 0, stream))
elif what == "ltv":
    n = 262_144
    eng.load_mlp(synth.make_mlp())
    x, rows = synth.make_ltv_inputs(n)
    d_x = torch.from_numpy(x.view(np.int16)).to(dev)
    d_rows = torch.from_numpy(rows.view(np.uint8).reshape(n, -1)).to(dev)
    d_o = torch.zeros((n, 32), dtype=torch.uint8, device=dev)
    d_raw = torch.zeros(n, dtype=torch.float32, device=dev)
    out["ms"] = timed(lambda: lib.igs_ltv_batch_device(eng.handle, d_x.data_ptr(), d_rows.data_ptr(), n, d_o.data_ptr(),
                                                       d_raw.data_ptr(), stream))
elif what in ("c3", "append"):
    H, A, Q = 10_000_000, 1_000_000, 1_000_000
    ev = synth.make_events(H, A)
    import time
    walls = []
    for _ in range(reps if what == "c3" else 1):
        t0 = time.perf_counter()
        eng.load_events(ev, A)
        walls.append((time.perf_counter() - t0) * 1e3)
    out["load_wall_ms"] = walls
    if what == "c3":
        q = synth.queries_of(synth.make_tx(Q, A, zipf_s=1.05))
        d_q = torch.from_numpy(q.view(np.uint8).reshape(Q, -1)).to(dev)
        d_v = torch.zeros((Q, 128), dtype=torch.uint8, device=dev)
        out["query_ms"] = timed(lambda: lib.igs_get_velocity_device(eng.handle, d_q.data_ptr(), Q, d_v.data_ptr(), stream))
    else:
        extra = synth.make_events(100_000 * (reps + 1), A, seed=77)
        order = np.argsort(extra.ts, kind="stable")      # arrival order = time order, later than the history
        shift = np.uint32(ev.ts.max() - extra.ts.min() + 1)
        app = []
        for k in range(reps + 1):
            sl = order[k * 100_000:(k + 1) * 100_000]
            part = pkg.Events(extra.account[sl], extra.ts[sl] + shift, extra.amount[sl], extra.tx_type[sl])
            t0 = time.perf_counter()
            eng.append_events(part)
            app.append((time.perf_counter() - t0) * 1e3)
        out["append_wall_ms"] = app
else:
    raise SystemExit(f"unknown scenario {what}")
torch.cuda.synchronize()
print(json.dumps(out), flush=True)
eng.close()
