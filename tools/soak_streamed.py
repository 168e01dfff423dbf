#!/usr/bin/env python
"""Soak of the streamed host-buffer path (igs_score_batch on large pinned batches: one kernel launch, the records copied
in behind it): many calls of varying ragged size that alternate between two different batches in the SAME pinned and
staging buffers, each compared with the staged path's results.  A producer warp that read a record before its chunk
arrived would see the other batch's record and the comparison would fail.

    python tools/soak_streamed.py [iterations, default 200]
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

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
n_players, n_events, n_max = 200_000, 2_000_000, 1_200_000
eng = pkg.Engine(0)
now = synth.T0 + 6 * 3600
eng.load_players(synth.make_players(n_players, seed=5, now=now))
eng.load_events(synth.make_events(n_events, n_players, seed=6, span=6 * 3600), n_players)
eng.load_gbdt(synth.make_gbdt(n_trees=60, seed=8))
cfg = pkg.default_config(model_kind=abi.MODEL_GBDT)
txs = [synth.make_tx(n_max, n_players, seed=70 + k, t_lo=synth.T0 + 3600, t_hi=now) for k in range(2)]
want = [eng.score_batch(cfg, t) for t in txs]          # pageable buffers: the staged path

p_tx = lib.igs_alloc_pinned(n_max * 32)
p_out = lib.igs_alloc_pinned(n_max * 8)
h_tx = np.frombuffer((C.c_uint8 * (n_max * 32)).from_address(p_tx), dtype=abi.tx_dtype)
h_out = np.frombuffer((C.c_uint8 * (n_max * 8)).from_address(p_out), dtype=abi.result_dtype)
rng = np.random.default_rng(3)
bad = 0
for it in range(iters):
    k = it & 1
    n = int(rng.integers(4 * 148 * 512, n_max + 1))
    h_tx[:n] = txs[k][:n]
    h_out.view(np.uint8)[:] = 0xEE
    st = lib.igs_score_batch(eng.handle, C.byref(cfg), h_tx.ctypes.data, n, h_out.ctypes.data, None, None)
    assert st == 0, lib.igs_last_error(eng.handle).decode()
    if h_out[:n].tobytes() != want[k][:n].tobytes():
        bad += 1
    assert (h_out.view(np.uint8)[n * 8:] == 0xEE).all(), "wrote past the batch"
print(json.dumps({"tool": "soak_streamed", "iterations": iters, "mismatching_calls": bad}), flush=True)
lib.igs_free_pinned(p_tx)
lib.igs_free_pinned(p_out)
eng.close()
sys.exit(1 if bad else 0)
