#!/usr/bin/env python
"""Cost of the identifier features on config C3 sizes: index build with device / IP
columns, the uniques query, the blacklist probe, and the fused path with them switched on.  Prints one
JSON object; run under `ncu --metrics gpu__time_duration.sum` for the per-kernel split."""
import json
import os
import sys
import time

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
import ctypes as C  # noqa: E402

H = int(os.environ.get("IGS_EVENTS", 10_000_000))
A = int(os.environ.get("IGS_ACCOUNTS", 1_000_000))
Q = int(os.environ.get("IGS_QUERIES", 1_000_000))

plain = synth.make_events(H, A)
ev = synth.make_event_ids(plain)
tx = synth.make_tx(Q, A, zipf_s=1.05)
q = synth.queries_of(tx)
eng = pkg.Engine(0)
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream().cuda_stream


def wall(fn, reps=3):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
    return best


def timed(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


build_plain = wall(lambda: eng.load_events(plain, A))
build_ids = wall(lambda: eng.load_events(ev, A))
extra = synth.make_event_ids(synth.make_events(600_000, A, seed=77), seed=78)
appends = []
for k in range(6):
    s = slice(k * 100_000, (k + 1) * 100_000)
    part = pkg.Events(extra.account[s], extra.ts[s], extra.amount[s], extra.tx_type[s], device=extra.device[s], ip=extra.ip[s])
    appends.append(wall(lambda: eng.append_events(part), reps=1))
eng.load_events(ev, A)

d_q = torch.from_numpy(q.view(np.uint8).reshape(Q, -1)).to(dev)
d_u = torch.zeros((Q, 8), dtype=torch.uint8, device=dev)
d_v = torch.zeros((Q, 128), dtype=torch.uint8, device=dev)
uniq_ms = timed(lambda: lib.igs_get_uniques_device(eng.handle, d_q.data_ptr(), Q, d_u.data_ptr(), stream))
vel_ms = timed(lambda: lib.igs_get_velocity_device(eng.handle, d_q.data_ptr(), Q, d_v.data_ptr(), stream))

# blacklists: 1M devices, 100k fingerprints, 1M IPs
rng = np.random.default_rng(3)
bl = [rng.integers(1, 1 << 63, n, dtype=np.uint64) for n in (1_000_000, 100_000, 1_000_000)]
load_ms = [wall(lambda k=k: eng.blacklist_load(k, bl[k]), reps=2) * 1e3 for k in range(3)]
ids = np.zeros(Q, dtype=abi.tx_ids_dtype)
for k, name in enumerate(("device", "fingerprint", "ip")):
    ids[name] = np.where(rng.random(Q) < 0.01, bl[k][rng.integers(0, len(bl[k]), Q)], rng.integers(1, 1 << 63, Q, dtype=np.uint64))
d_ids = torch.from_numpy(ids.view(np.uint8).reshape(Q, -1)).to(dev)
d_hit = torch.zeros(Q, dtype=torch.uint8, device=dev)
chk_ms = timed(lambda: lib.igs_blacklist_check_device(eng.handle, d_ids.data_ptr(), Q, d_hit.data_ptr(), stream))
hit_rate = float(d_hit.float().mean().item())

# fused path: with and without the identifier inputs
model = synth.make_gbdt(n_trees=500)
eng.load_gbdt(model)
eng.load_players(synth.make_players(A))
cfg = pkg.default_config(model_kind=abi.MODEL_GBDT)
d_tx = torch.from_numpy(tx.view(np.uint8).reshape(Q, -1)).to(dev)
d_out = torch.zeros((Q, 8), dtype=torch.uint8, device=dev)
fused_ids_ms = timed(lambda: lib.igs_score_batch_ids_device(eng.handle, C.byref(cfg), d_tx.data_ptr(), d_ids.data_ptr(), Q,
                                                            d_out.data_ptr(), None, None, stream), reps=10)
fused_uniq_ms = timed(lambda: lib.igs_score_batch_device(eng.handle, C.byref(cfg), d_tx.data_ptr(), Q, d_out.data_ptr(),
                                                         None, None, stream), reps=10)
eng.load_events(plain, A)
fused_plain_ms = timed(lambda: lib.igs_score_batch_device(eng.handle, C.byref(cfg), d_tx.data_ptr(), Q, d_out.data_ptr(),
                                                          None, None, stream), reps=10)
print(json.dumps({
    "events": H, "accounts": A, "queries": Q,
    "index_build_wall_ms": {"plain": build_plain * 1e3, "with_identifiers": build_ids * 1e3,
                            "note": "igs_load_events(_ids) end to end incl. H2D of 17 / 33 B per event"},
    "append_100k_with_identifiers_ms": {"first": appends[0] * 1e3, "steady_median": float(np.median(appends[1:])) * 1e3},
    "uniques_query_ms": uniq_ms, "velocity_query_ms": vel_ms,
    "blacklist": {"load_ms": load_ms, "sizes": [len(b) for b in bl], "check_ms_per_1M": chk_ms, "hit_rate": hit_rate,
                  "probes_gbs": Q * 24 / (chk_ms * 1e-3) / 1e9},
    "fused_ms_per_1M": {"plain_history": fused_plain_ms, "uniques_from_index": fused_uniq_ms,
                        "uniques_and_blacklist": fused_ids_ms},
}), flush=True)
eng.close()
