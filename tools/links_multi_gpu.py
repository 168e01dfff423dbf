#!/usr/bin/env python
"""Linked accounts across real GPUs: one process per GPU, NCCL all-to-all over NVLink.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      tools/links_multi_gpu.py [--events-per-shard E] [--accounts A] [--check]

Every rank builds the same synthetic global history (seeded), keeps the events of its own accounts
(shard = igs_shard_of(account)), loads them, and runs links.build_links.  --check compares every shard's
table with the CPU oracle run on the global history (small sizes only).  Rank 0 prints one JSON line with
the stage record counts and the time of the whole job (max over ranks, CUDA events around it)."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--events-per-shard", type=int, default=2_000_000)
    ap.add_argument("--accounts", type=int, default=1_000_000, help="global account count")
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--peer", action="store_true",
                    help="exchange through peer-memory windows (stores + atomics over NVLink from inside the "
                         "producing kernels) instead of NCCL all-to-all")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ge.build_library()
    pkg = ge.load_package()
    abi, links = pkg.abi, pkg.links
    lib = abi.load_library()
    from tests import synth
    from tests.test_links_gpu import make_history

    n_ev = args.events_per_shard * world
    ev = make_history(pkg, n_ev, args.accounts, seed=2024, shared_rate=0.05, pools=max(args.accounts // 20, 10))
    # placement: vectorised igs_shard_of (splitmix64(account) % world)
    g = np.arange(args.accounts, dtype=np.uint64)
    with np.errstate(over="ignore"):
        shard_of = (synth.mix64(g + np.uint64(0x9E3779B97F4A7C15)) % np.uint64(world)).astype(np.uint32)
    assert int(shard_of[12345 % args.accounts]) == lib.igs_shard_of(12345 % args.accounts, 0, world)
    dense = np.zeros(args.accounts, dtype=np.uint32)
    mine = np.nonzero(shard_of == rank)[0]
    for s in range(world):
        idx = np.nonzero(shard_of == s)[0]
        dense[idx] = np.arange(idx.size)
    m = shard_of[ev.account] == rank
    eng = pkg.Engine(local)
    eng.load_events(pkg.Events(dense[ev.account[m]], ev.ts[m], ev.amount[m], ev.tx_type[m], device=ev.device[m],
                               ip=ev.ip[m]), mine.size)
    if args.peer:
        # capacity: every event could carry a device owned by one shard; links bounded generously
        links.setup_peer_windows(eng, rank, world, cap_pairs=n_ev, cap_links=4 * n_ev)
    times, stats = [], {"pairs_sent": 0, "links_sent": 0}
    for _ in range(args.reps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if args.peer:
            links.build_links_peer(eng, rank, world)
        else:
            stats = links.build_links(eng, rank, world)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times.append(float(t.item()))
    ok = True
    if args.check:
        from oracle import oracle as orc
        gid = (shard_of.astype(np.uint64) << np.uint64(32)) | dense.astype(np.uint64)
        want = orc.links(ev.device, gid[ev.account], gid[mine])
        got = eng.get_links(np.arange(mine.size, dtype=np.uint32))
        ok = got.tobytes() == want.tobytes()
        crossing = int(((got["linked"][:, 0] >> np.uint64(32)) != rank)[got["n_linked"] > 0].sum())
        flag = torch.tensor([int(not ok), crossing], device="cuda")      # failures and shard-crossing links, summed
        if world > 1:
            dist.all_reduce(flag)
        ok = int(flag[0].item()) == 0 and (world == 1 or int(flag[1].item()) > 0)
    tot = torch.tensor([stats["pairs_sent"], stats["links_sent"], int(m.sum())], device="cuda", dtype=torch.int64)
    if world > 1:
        dist.all_reduce(tot)
    if rank == 0:
        best = min(times)
        print(json.dumps({"job": "linked accounts (igs_links_pairs -> all-to-all -> igs_links_group -> all-to-all -> "
                                 "igs_links_finish)", "exchange": "peer-memory windows (fused into the kernels)" if args.peer
                          else "NCCL all_to_all_single", "n_gpus": world, "events": int(tot[2].item()),
                          "accounts": args.accounts, "pairs_exchanged": int(tot[0].item()),
                          "links_exchanged": int(tot[1].item()), "ms_best": best, "ms_all": times,
                          "events_per_s": int(tot[2].item()) / (best * 1e-3),
                          "exchange_bytes": 16 * (int(tot[0].item()) + int(tot[1].item())),
                          "checked_against_oracle": bool(args.check), "ok": ok}), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
