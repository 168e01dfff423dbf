#!/usr/bin/env python
"""K1 shard partition alone: 12.5M tx (the 8-GPU C5 slice) into 8 shards on one GPU, CUDA-event time per call and the
fraction of the measured HBM peak on 68 algorithmic bytes per tx (32 read + 32 written + 4 of original index).

    python tools/bench_partition.py [n_tx] [n_shards]
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

ge.build_library()
pkg = ge.load_package()
lib = pkg.abi.load_library()
from tests import synth  # noqa: E402
import torch  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 12_500_000
shards = int(sys.argv[2]) if len(sys.argv) > 2 else 8
dev = torch.device("cuda", 0)
stream = torch.cuda.current_stream().cuda_stream
eng = pkg.Engine(0)
base_n = 2_000_000
inc = synth.make_tx(base_n, 4_000_000, seed=7)
d_in = torch.from_numpy(inc.view(np.uint8).reshape(base_n, -1)).to(dev).repeat(-(-n // base_n), 1)[:n].contiguous()
d_out = torch.empty_like(d_in)
d_perm = torch.zeros(n, dtype=torch.int32, device=dev)
d_off = torch.zeros(shards + 1, dtype=torch.int32, device=dev)
ms = []
for it in range(8):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    rc = lib.igs_partition_tx_device(eng.handle, d_in.data_ptr(), n, shards, d_out.data_ptr(), d_perm.data_ptr(),
                                     d_off.data_ptr(), stream)
    b.record()
    torch.cuda.synchronize()
    assert rc == 0, rc
    ms.append(a.elapsed_time(b))
# the partition is stable: per shard, original indices ascend; and every record landed in its shard
perm = d_perm.cpu().numpy()
off = d_off.cpu().numpy()
acc = d_out.cpu().numpy().view(inc.dtype).reshape(-1)["account"]
want = pkg.sharding.shard_of(acc, shards)
for s in range(shards):
    seg = slice(off[s], off[s + 1])
    assert (want[seg] == s).all()
    assert (np.diff(perm[seg].astype(np.int64)) > 0).all()
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
best = float(np.median(ms[2:]))
print(json.dumps({"n": n, "shards": shards, "ms": ms, "median_ms": best, "gbs": n * 68 / best / 1e6,
                  "frac_of_hbm_peak": n * 68 / best / 1e6 / peak}), flush=True)
eng.close()
