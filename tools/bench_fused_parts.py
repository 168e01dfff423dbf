#!/usr/bin/env python
"""Where does the fused path's extra time go?  Times igs_score_batch_device on 1M tx with
(a) players + event index, (b) players only, (c) nothing loaded, against path A on the same GPU."""
import ctypes as C, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
ge.build_library()
pkg = ge.load_package(); abi = pkg.abi; lib = abi.load_library()
from tests import synth
import torch
n, n_pl, n_ev = 1_000_000, 500_000, 5_000_000
dev = torch.device("cuda", 0); stream = torch.cuda.current_stream().cuda_stream
cfg = pkg.default_config(model_kind=abi.MODEL_GBDT)
model = synth.make_gbdt(500)
tx = synth.make_tx(n, n_pl)
d_tx = torch.from_numpy(tx.view(np.uint8).reshape(n, -1)).to(dev)
d_out = torch.zeros((n, 8), dtype=torch.uint8, device=dev)
def timed(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps
res = {}
for label, load_pl, load_ev in (("nothing_loaded", False, False), ("players_only", True, False), ("players_and_events", True, True)):
    eng = pkg.Engine(0); eng.load_gbdt(model)
    if load_pl: eng.load_players(synth.make_players(n_pl))
    if load_ev: eng.load_events(synth.make_events(n_ev, n_pl), n_pl)
    res[label] = timed(lambda: lib.igs_score_batch_device(eng.handle, C.byref(cfg), d_tx.data_ptr(), n, d_out.data_ptr(), None, None, stream))
    if load_ev:      # the same batch ordered by account: what an account-sorted batch would buy (locality of the index gathers)
        d_sorted = torch.from_numpy(tx[np.argsort(tx["account"], kind="stable")].view(np.uint8).reshape(n, -1)).to(dev)
        res["players_and_events_batch_sorted_by_account"] = timed(lambda: lib.igs_score_batch_device(
            eng.handle, C.byref(cfg), d_sorted.data_ptr(), n, d_out.data_ptr(), None, None, stream))
        res["distinct_accounts_in_batch"] = int(np.unique(tx["account"]).size)
    eng.close()
print(json.dumps(res))
