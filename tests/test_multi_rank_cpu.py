"""CPU, world_size 2 over gloo: the N>1 path's host logic.  Each rank takes its shard of one mixed tx
batch (same rule as the device partition), scores it independently — here with the oracle standing in
for the GPU, as the checker — and only results are gathered; the merged result must equal scoring the
whole batch in one place (shards are independent: no feature crosses an account)."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, pickle
import numpy as np
import torch.distributed as dist
sys.path.insert(0, os.environ["IGS_ROOT"])
import __graft_entry__ as ge
pkg = ge.load_package()
from oracle import oracle as orc
from tests import synth

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
abi, sh = pkg.abi, pkg.sharding
A, H, N = 4000, 30000, 20000
players = synth.make_players(A, seed=1, now=synth.T0 + 6 * 3600)
ev = synth.make_events(H, A, seed=2, span=6 * 3600)
tx = synth.make_tx(N, A, seed=3, t_lo=synth.T0 + 3600, t_hi=synth.T0 + 6 * 3600)
cfg = orc.default_config(model_kind=abi.MODEL_MOCK)

# every rank derives the same partition; rank r keeps shard r
acc_shard, acc_local = sh.local_index(np.arange(A, dtype=np.uint32), world)
shards, perm, offsets = sh.split_by_shard(tx, world)
mine = shards[rank].copy()
mine["account"] = acc_local[mine["account"]]                      # global id -> dense index on this shard
my_players = players[acc_shard == rank]
em = acc_shard[ev.account] == rank
my_ev = pkg.Events(acc_local[ev.account[em]], ev.ts[em], ev.amount[em], ev.tx_type[em])
res = orc.score_batch(cfg, None, my_players, my_ev, mine)

gathered = [None] * world
dist.all_gather_object(gathered, res.tobytes())                    # results only: 8 B per tx
if rank == 0:
    per = [np.frombuffer(b, dtype=abi.result_dtype) for b in gathered]
    merged = sh.merge_results(per, perm)
    whole = orc.score_batch(cfg, None, players, ev, tx)
    ok = merged.tobytes() == whole.tobytes()
    sizes = [int(offsets[s + 1] - offsets[s]) for s in range(world)]
    print("RESULT " + str(ok) + " " + str(sizes), flush=True)
dist.barrier()
dist.destroy_process_group()
'''


def test_two_ranks_score_their_shards_independently(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, IGS_ROOT=ROOT, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29517", str(script)],
                       capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT")]
    assert line and line[0].split()[1] == "True", r.stdout[-2000:]
    sizes = json.loads(line[0].split(" ", 2)[2])
    assert sum(sizes) == 20000 and min(sizes) > 7000        # hash shards are balanced


def test_split_and_merge_roundtrip(pkg):
    from tests import synth
    sh = pkg.sharding
    tx = synth.make_tx(5000, 300, seed=9)
    for n in (1, 2, 3, 8):
        shards, perm, offsets = sh.split_by_shard(tx, n)
        assert [s.shape[0] for s in shards] == [int(offsets[i + 1] - offsets[i]) for i in range(n)]
        for s_idx, s in enumerate(shards):
            assert (sh.shard_of(s["account"], n) == s_idx).all()
        assert sh.merge_results(shards, perm).tobytes() == tx.tobytes()


def test_reference_arm_under_torchrun_prints_one_line():
    """bench.py --impl reference under torchrun: rank 0 alone runs and prints; the others exit 0."""
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29518", os.path.join(ROOT, "bench.py"),
                        "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["value"] > 0 and j["cpu_baseline"]["kind"] == "port"
    assert j["e2e"]["h2d_bytes_per_step"] == 0 and j["n_gpus"] == 2


LINKS_WORKER = r'''
import os, sys
import numpy as np
import torch.distributed as dist
sys.path.insert(0, os.environ["IGS_ROOT"])
import __graft_entry__ as ge
pkg = ge.load_package()
from oracle import oracle as orc
from tests import synth
from tests.test_links_gpu import make_history

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
abi, links, sh = pkg.abi, pkg.links, pkg.sharding
K = abi.LINK_MAX_ACCOUNTS
A, H = 1500, 20000
ev = make_history(pkg, H, A, seed=7, shared_rate=0.1, pools=150)
acc_shard, acc_local = sh.local_index(np.arange(A, dtype=np.uint32), world)
gid = links.global_account_id(0, 0) * 0 + ((acc_shard.astype(np.uint64) << np.uint64(32)) | acc_local.astype(np.uint64))
mine = acc_shard[ev.account] == rank


class HostEngine:
    """Stand-in for igs.Engine on host arrays: the three stages with the library's output convention
    (records grouped by destination shard, per-shard counts).  Only the driver and the exchange are under
    test here; the device stages are tested on the GPU."""
    device = 0

    def __init__(self):
        self.rows = None

    @staticmethod
    def grouped(rec, dest):
        order = np.argsort(dest, kind="stable")
        return rec[order], np.bincount(dest, minlength=world).astype(np.uint64)

    def links_pairs(self, shard, n_shards):
        d, a = ev.device[mine], gid[ev.account[mine]]
        keep = d != 0
        rec = np.zeros(int(keep.sum()), dtype=abi.link_pair_dtype)
        rec["device"], rec["account"] = d[keep], a[keep]
        return self.grouped(rec, sh.shard_of(rec["device"], n_shards))

    def links_group(self, pairs, n, n_shards):
        pairs = np.unique(pairs[:n])
        out = []
        for dev in np.unique(pairs["device"]):
            accts = pairs["account"][pairs["device"] == dev]
            if 2 <= accts.size <= K:
                for x in accts:
                    out += [(x, y) for y in accts if y != x]
        rec = np.array(out, dtype=abi.link_dtype) if out else np.zeros(0, dtype=abi.link_dtype)
        return self.grouped(rec, (rec["account"] >> np.uint64(32)).astype(np.int64))

    def links_finish(self, lk, n, shard):
        lk = np.unique(lk[:n])
        n_rows = int((acc_shard == shard).sum())
        rows = np.zeros(n_rows, dtype=abi.links_row_dtype)
        for i in range(n_rows):
            others = np.sort(lk["other"][lk["account"] == links.global_account_id(shard, i)])
            rows["n_linked"][i] = others.size
            rows["linked"][i, :min(3, others.size)] = others[:3]
        self.rows = rows


def exchange(rec, counts):
    """all-to-all over gloo: rank r's group s goes to rank s; delivery is rank-major."""
    starts = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    send = [rec[starts[s]:starts[s + 1]].tobytes() for s in range(world)]
    everyone = [None] * world
    dist.all_gather_object(everyone, send)
    got = np.frombuffer(b"".join(everyone[src][rank] for src in range(world)), dtype=rec.dtype)
    return None, got, got.shape[0]


eng = HostEngine()
stats = links.build_links(eng, rank, world, exchange=exchange)
idx = np.nonzero(acc_shard == rank)[0]
want = orc.links(ev.device, gid[ev.account], gid[idx])
ok = eng.rows.tobytes() == want.tobytes()
crossing = int(((want["linked"][:, 0] >> np.uint64(32)) != rank)[want["n_linked"] > 0].sum())
res = [None] * world
dist.all_gather_object(res, (ok, crossing, stats["pairs_received"], stats["links_received"]))
if rank == 0:
    print("RESULT " + str(all(r[0] for r in res)) + " " + str(sum(r[1] for r in res)) + " " + str(res), flush=True)
dist.barrier()
dist.destroy_process_group()
'''


def test_two_ranks_build_links_through_the_exchange(tmp_path):
    """links.build_links under world_size 2 (gloo): the driver's stage order, the grouped-by-shard buffer
    convention and the rank-major delivery of the all-to-all, with a host stand-in for the device stages;
    each rank's table must equal the oracle's on the global history, and links must cross the shards."""
    script = tmp_path / "links_worker.py"
    script.write_text(LINKS_WORKER)
    env = dict(os.environ, IGS_ROOT=ROOT, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29519", str(script)],
                       capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT")]
    assert line and line[0].split()[1] == "True", r.stdout[-2000:]
    assert int(line[0].split()[2]) > 0, "no link crossed a shard boundary"
