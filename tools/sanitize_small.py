#!/usr/bin/env python
"""Tiny invocation of every device entry point, meant to run under
`compute-sanitizer --tool memcheck` (one tool per gpurun call, see B200_PROFILING.md)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

ge.build_library()
pkg = ge.load_package()
abi = pkg.abi
from tests import synth  # noqa: E402

eng = pkg.Engine(0)
# path A: ragged tile, chunk tail, leaves
model = synth.make_gbdt(n_trees=47, seed=1)
eng.load_gbdt(model)
feats, rows = synth.make_feature_batch(1000, seed=2, invalid_rate=0.1)
for kind in (abi.MODEL_NONE, abi.MODEL_MOCK, abi.MODEL_GBDT):
    out = eng.score_features(pkg.default_config(model_kind=kind), feats, rows)
out, leaves = eng.score_features(pkg.default_config(model_kind=abi.MODEL_GBDT), feats, rows, want_leaves=True)
# velocity + fused path
players = synth.make_players(300, seed=3)
ev = synth.make_events(5000, 300, seed=4, span=3600)
tx = synth.make_tx(1500, 300, seed=5, t_lo=synth.T0, t_hi=synth.T0 + 3600)
tx["account"][::50] = 400
eng.load_players(players)
eng.load_events(ev, 300)
v = eng.get_velocity(synth.queries_of(tx))
r, f, vv = eng.score_batch(pkg.default_config(model_kind=abi.MODEL_GBDT), tx, want_features=True, want_velocity=True)
# pipelined host path (> 2 chunks)
feats, rows = synth.make_feature_batch(300_000, seed=6)
out = eng.score_features(pkg.default_config(model_kind=abi.MODEL_MOCK), feats, rows)
# partition (both entry points) through torch device buffers
import torch  # noqa: E402
n = 5000
txp = synth.make_tx(n, 1000, seed=7)
dev = torch.device("cuda", 0)
d_tx = torch.from_numpy(txp.view(np.uint8).reshape(n, 32).copy()).to(dev)
d_out = torch.zeros((n, 32), dtype=torch.uint8, device=dev)
d_perm = torch.zeros(n, dtype=torch.int32, device=dev)
d_off = torch.zeros(9, dtype=torch.int32, device=dev)
for call in (eng.partition_tx_device, eng.partition_tx_static):
    call(d_tx.data_ptr(), n, 8, d_out.data_ptr(), d_perm.data_ptr(), d_off.data_ptr(), torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
# LTV
eng.load_mlp(synth.make_mlp(seed=8))
x, lrows = synth.make_ltv_inputs(300, seed=9)
o, raw = eng.ltv_batch(x, lrows, want_raw=True)
eng.close()
print("sanitize_small: all entry points ran;", int(out["score"].sum()), int(v["n_prior"].sum()), float(raw.sum()))
