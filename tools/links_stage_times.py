#!/usr/bin/env python
"""Per-stage wall time of the linked-accounts job on one GPU (no exchange), steady state."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
ge.build_library()
pkg = ge.load_package()
from tests.test_links_gpu import make_history
import torch
n_ev, n_acc = int(os.environ.get("IGS_EVENTS", 5_000_000)), int(os.environ.get("IGS_ACCOUNTS", 500_000))
ev = make_history(pkg, n_ev, n_acc, seed=2024, shared_rate=0.05, pools=n_acc // 20)
eng = pkg.Engine(0)
eng.load_events(ev, n_acc)
rows = []
for rep in range(4):
    t = [time.perf_counter()]
    p, c = eng.links_pairs(0, 1); torch.cuda.synchronize(); t.append(time.perf_counter())
    l, lc = eng.links_group(p, int(c.sum()), 1); torch.cuda.synchronize(); t.append(time.perf_counter())
    eng.links_finish(l, int(lc.sum()), 0); torch.cuda.synchronize(); t.append(time.perf_counter())
    rows.append([round((b - a) * 1e3, 3) for a, b in zip(t, t[1:])])
print(json.dumps({"events": n_ev, "pairs": int(c.sum()), "links": int(lc.sum()), "ms_pairs_group_finish": rows}))
