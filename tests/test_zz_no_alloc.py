"""GPU: igs_partition_tx_static, the partition entry point that allocates nothing.

It exists for callers that cannot afford an allocator call on this path: a shard's ingest thread
that partitions batches while scoring calls are in flight on the same GPU (cudaMalloc / cudaFree
stall the whole device), and deployments with a GPU memory budget fixed at start-up.  Its counters
live in a 1 MiB array inside the library, shared by every call on the device and never cleared, so
this test makes two calls in a row with different inputs — the second smaller than the first, so it
runs on top of the first call's leftovers — and requires both to reproduce, byte for byte, what the
allocating entry point igs_partition_tx_device returns for the same input (integer work: no
tolerance).  Collected last on purpose; it must pass on its own and after every other test."""
import numpy as np
import pytest

from tests import synth
from tests.test_partition_gpu import run_partition

pytestmark = pytest.mark.gpu


def test_static_partition_matches_allocating_entry_point_twice(engine, pkg):
    for n, n_shards, seed in ((300_000, 8, 1), (70_001, 5, 2)):
        tx = synth.make_tx(n, 50_000, seed=seed)
        want_out, want_perm, want_off = run_partition(engine, pkg, tx, n_shards)
        got_out, got_perm, got_off = run_partition(engine, pkg, tx, n_shards, call=engine.partition_tx_static)
        assert np.array_equal(got_off, want_off), f"offsets differ (n={n})"
        assert np.array_equal(got_perm, want_perm), f"permutation differs (n={n})"
        assert got_out.tobytes() == want_out.tobytes(), f"records differ (n={n})"
        assert int(got_off[-1]) == n


def test_static_partition_edge_sizes_and_limit(engine, pkg):
    abi = pkg.abi
    # empty batch: offsets all zero, nothing launched
    _, _, off = run_partition(engine, pkg, synth.make_tx(0, 10, seed=3), 4, call=engine.partition_tx_static)
    assert off.tolist() == [0] * 5
    # largest request the built-in array serves with 256 shards
    n_ok = abi.PARTITION_STATIC_WORDS // 256 * 2048
    tx = synth.make_tx(n_ok, 1000, seed=4)
    got = run_partition(engine, pkg, tx, 256, call=engine.partition_tx_static)
    want = run_partition(engine, pkg, tx, 256)
    assert got[0].tobytes() == want[0].tobytes() and np.array_equal(got[2], want[2])
    # one block more is refused with a message that names the limit; nothing runs
    with pytest.raises(pkg.IgsError) as ei:
        engine.partition_tx_static(1, n_ok + 1, 256, 1, 0, 1)
    assert ei.value.status == abi.ERR_INVALID
    assert str(abi.PARTITION_STATIC_WORDS) in str(ei.value) and "igs_partition_tx_device" in str(ei.value)
