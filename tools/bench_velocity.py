#!/usr/bin/env python
"""Kernel-level timing of the HBM-bound side of the path on config C3 (10M events, 1M accounts, 1M
queries): index build (K2 sort + K3 scans), velocity query, shard partition (K1).  Prints one JSON
object; run it under `ncu --metrics gpu__time_duration.sum` for the per-kernel split."""
import ctypes as C
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

H = int(os.environ.get("IGS_EVENTS", 10_000_000))
A = int(os.environ.get("IGS_ACCOUNTS", 1_000_000))
Q = int(os.environ.get("IGS_QUERIES", 1_000_000))
PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0

ev = synth.make_events(H, A)
tx = synth.make_tx(Q, A, zipf_s=1.05)
q = synth.queries_of(tx)
eng = pkg.Engine(0)
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream().cuda_stream

# index build: wall clock of igs_load_events (H2D of 17 B/event + sort + scans), best of 3
build = []
for _ in range(3):
    t0 = time.perf_counter()
    eng.load_events(ev, A)
    build.append(time.perf_counter() - t0)

# the same history arriving in time order (a log / a queue): stable sort by account only
order = np.argsort(ev.ts, kind="stable")
ev_sorted = pkg.Events(ev.account[order], ev.ts[order], ev.amount[order], ev.tx_type[order])
build_sorted = []
for _ in range(3):
    t0 = time.perf_counter()
    eng.load_events(ev_sorted, A)
    build_sorted.append(time.perf_counter() - t0)
eng.load_events(ev, A)

# incremental ingest: append 100k-event micro-batches to the resident 10M-event history
extra = synth.make_events(600_000, A, seed=77)
order = np.argsort(extra.ts, kind="stable")            # a live feed: in time order, after the resident history
shift = np.uint32(ev.ts.max() - extra.ts.min() + 1)
extra = pkg.Events(extra.account[order], extra.ts[order] + shift, extra.amount[order], extra.tx_type[order])
appends = []
for k in range(6):
    part = pkg.Events(extra.account[k * 100_000:(k + 1) * 100_000], extra.ts[k * 100_000:(k + 1) * 100_000],
                      extra.amount[k * 100_000:(k + 1) * 100_000], extra.tx_type[k * 100_000:(k + 1) * 100_000])
    t0 = time.perf_counter()
    eng.append_events(part)
    appends.append(time.perf_counter() - t0)

d_q = torch.from_numpy(q.view(np.uint8).reshape(Q, -1)).to(dev)
d_v = torch.zeros((Q, 128), dtype=torch.uint8, device=dev)


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


vel_ms = timed(lambda: lib.igs_get_velocity_device(eng.handle, d_q.data_ptr(), Q, d_v.data_ptr(), stream))

N = 10_000_000
big = synth.make_tx(N, A, seed=5)
d_tx = torch.from_numpy(big.view(np.uint8).reshape(N, -1)).to(dev)
d_out = torch.zeros((N, 32), dtype=torch.uint8, device=dev)
d_perm = torch.zeros(N, dtype=torch.int32, device=dev)
d_off = torch.zeros(9, dtype=torch.int32, device=dev)
part_ms = timed(lambda: lib.igs_partition_tx_device(eng.handle, d_tx.data_ptr(), N, 8, d_out.data_ptr(),
                                                    d_perm.data_ptr(), d_off.data_ptr(), stream))

vel_bytes = Q * (16 + 128)                       # query record in + velocity block out (gathers excluded)
vel_stream_bytes = (H + Q) * 17 + Q * 100        # SURVEY §8d streaming formulation
part_bytes = N * (32 + 32 + 4)
print(json.dumps({
    "events": H, "accounts": A, "queries": Q, "hbm_peak_gbs": PEAK,
    "index_build": {"wall_s_best": min(build), "events_per_s": H / min(build),
                    "note": "igs_load_events end to end: H2D of 17 B/event, radix sort, gathers, scans"},
    "index_build_time_ordered_arrival": {"wall_s_best": min(build_sorted), "events_per_s": H / min(build_sorted)},
    "append_100k_events": {"wall_ms_first": appends[0] * 1e3, "wall_ms_steady_median": float(np.median(appends[1:])) * 1e3,
                           "note": "igs_append_events onto the resident history: H2D of the batch + rebuild of the delta index"},
    "velocity_query": {"ms": vel_ms, "queries_per_s": Q / (vel_ms * 1e-3),
                       "gbs_query_io": vel_bytes / (vel_ms * 1e-3) / 1e9,
                       "gbs_vs_streaming_formulation": vel_stream_bytes / (vel_ms * 1e-3) / 1e9,
                       "frac_of_peak_streaming_formulation": vel_stream_bytes / (vel_ms * 1e-3) / 1e9 / PEAK},
    "partition_8_shards": {"n": N, "ms": part_ms, "tx_per_s": N / (part_ms * 1e-3),
                           "gbs": part_bytes / (part_ms * 1e-3) / 1e9,
                           "frac_of_peak": part_bytes / (part_ms * 1e-3) / 1e9 / PEAK},
}), flush=True)
eng.close()
