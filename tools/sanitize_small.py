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
# identifiers: uniques in the index, blacklist sets, fused path with ids; linked accounts (both exchanges)
evi = synth.make_event_ids(ev, seed=10)
eng.load_events(evi, 300)
eng.append_events(synth.make_event_ids(synth.make_events(700, 300, seed=11, span=3600), seed=12))
u = eng.get_uniques(synth.queries_of(tx))
for kind in (abi.BL_DEVICE, abi.BL_FINGERPRINT, abi.BL_IP):
    eng.blacklist_load(kind, evi.device[:50] if kind == abi.BL_DEVICE else np.arange(1, 40, dtype=np.uint64))
    eng.blacklist_add(kind, np.arange(100, 3000, dtype=np.uint64))
ids = np.zeros(tx.shape[0], dtype=abi.tx_ids_dtype)
ids["device"], ids["fingerprint"], ids["ip"] = evi.device[:tx.shape[0]], 7, evi.ip[:tx.shape[0]]
hit = eng.blacklist_check(ids)
r2 = eng.score_batch(pkg.default_config(model_kind=abi.MODEL_GBDT), tx, ids=ids)
pkg.links.build_links(eng)
rows0 = eng.get_links(np.arange(310, dtype=np.uint32))
pkg.links.setup_peer_windows(eng, 0, 1, cap_pairs=6000, cap_links=200_000)
pkg.links.build_links_peer(eng)
assert eng.get_links(np.arange(310, dtype=np.uint32)).tobytes() == rows0.tobytes()
ab = eng.check_bonus_abuse(np.arange(310, dtype=np.uint32))
for n0 in (0, 1):                                   # degenerate sizes
    eng.get_uniques(synth.queries_of(tx[:n0])); eng.blacklist_check(ids[:n0]); eng.get_links(np.arange(n0, dtype=np.uint32))
    eng.score_batch(pkg.default_config(model_kind=abi.MODEL_MOCK), tx[:n0], ids=ids[:n0])
eng.load_events(pkg.Events(np.zeros(0, np.uint32), np.zeros(0, np.uint32), np.zeros(0, np.int64), np.zeros(0, np.uint8),
                           device=np.zeros(0, np.uint64), ip=np.zeros(0, np.uint64)), 5)
pkg.links.build_links(eng)                          # empty history
pkg.links.build_links_peer(eng)
eng.close()
print("sanitize_small: all entry points ran;", int(u["devices_24h"].sum()), int(hit.sum()), int(rows0["n_linked"].sum()),
      int(ab["is_abuser"].sum()), int(out["score"].sum()), int(v["n_prior"].sum()), float(raw.sum()))
