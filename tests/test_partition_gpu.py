"""GPU: shard partition (K1) through the C ABI vs a numpy restatement of the same rule
(stable grouping by splitmix64(account) % n_shards)."""
import numpy as np
import pytest

from tests import synth

pytestmark = pytest.mark.gpu


def shard_of(account: np.ndarray, n_shards: int) -> np.ndarray:
    with np.errstate(over="ignore"):
        x = account.astype(np.uint64) + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
    return (x % np.uint64(n_shards)).astype(np.uint32)


def expected_partition(tx, n_shards):
    sh = shard_of(tx["account"], n_shards)
    perm = np.argsort(sh, kind="stable").astype(np.uint32)
    offsets = np.concatenate([[0], np.cumsum(np.bincount(sh, minlength=n_shards))]).astype(np.uint32)
    return tx[perm], perm, offsets


def run_partition(engine, pkg, tx, n_shards, call=None):
    import torch
    abi = pkg.abi
    n = tx.shape[0]
    dev = torch.device("cuda", 0)
    d_tx = torch.from_numpy(tx.view(np.uint8).reshape(n, 32).copy()).to(dev) if n else torch.zeros((1, 32), dtype=torch.uint8, device=dev)
    d_out = torch.full((max(n, 1), 32), 0xEE, dtype=torch.uint8, device=dev)
    d_perm = torch.full((max(n, 1),), -1, dtype=torch.int32, device=dev)
    d_off = torch.full((n_shards + 1,), -1, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    (call or engine.partition_tx_device)(d_tx.data_ptr(), n, n_shards, d_out.data_ptr(), d_perm.data_ptr(),
                                         d_off.data_ptr(), stream)
    torch.cuda.synchronize()
    out = np.frombuffer(d_out.cpu().numpy().tobytes(), dtype=abi.tx_dtype)[:n]
    perm = d_perm.cpu().numpy().view(np.uint32)[:n]
    off = d_off.cpu().numpy().view(np.uint32)
    return out, perm, off


@pytest.mark.parametrize("n,n_shards", [(0, 8), (1, 1), (1, 8), (2047, 2), (2048, 8), (2049, 8), (50000, 4),
                                         (50000, 256), (300000, 8), (1_000_000, 8)])
def test_partition_matches_stable_grouping(engine, pkg, n, n_shards):
    tx = synth.make_tx(n, 100_000, seed=n + n_shards)
    out, perm, off = run_partition(engine, pkg, tx, n_shards)
    want_out, want_perm, want_off = expected_partition(tx, n_shards)
    assert np.array_equal(off, want_off)
    assert np.array_equal(perm, want_perm)
    assert out.tobytes() == want_out.tobytes()


def test_partition_of_records_that_are_only_16_byte_aligned(engine, pkg):
    """Record arrays at a 16-byte offset of their allocation take the kernel variant without 256-bit moves."""
    import torch
    abi = pkg.abi
    n, n_shards = 70_001, 8
    tx = synth.make_tx(n, 100_000, seed=77)
    dev = torch.device("cuda", 0)
    raw = torch.zeros(n * 32 + 16, dtype=torch.uint8, device=dev)
    raw[16:] = torch.from_numpy(tx.view(np.uint8).reshape(-1).copy()).to(dev)
    out_raw = torch.zeros(n * 32 + 16, dtype=torch.uint8, device=dev)
    d_perm = torch.full((n,), -1, dtype=torch.int32, device=dev)
    d_off = torch.full((n_shards + 1,), -1, dtype=torch.int32, device=dev)
    assert (raw.data_ptr() + 16) % 32 == 16
    engine.partition_tx_device(raw.data_ptr() + 16, n, n_shards, out_raw.data_ptr() + 16, d_perm.data_ptr(),
                               d_off.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    out = np.frombuffer(out_raw[16:].cpu().numpy().tobytes(), dtype=abi.tx_dtype)
    want_out, want_perm, want_off = expected_partition(tx, n_shards)
    assert np.array_equal(d_off.cpu().numpy().view(np.uint32), want_off)
    assert np.array_equal(d_perm.cpu().numpy().view(np.uint32), want_perm)
    assert out.tobytes() == want_out.tobytes()


def test_partition_agrees_with_host_helper(engine, pkg):
    lib = pkg.abi.load_library()
    acc = np.arange(5000, dtype=np.uint32) * 7919
    want = np.array([lib.igs_shard_of(int(a), 0, 8) for a in acc], np.uint32)
    assert np.array_equal(shard_of(acc, 8), want)


def test_partition_rejects_bad_shard_count(engine, pkg):
    with pytest.raises(pkg.IgsError) as ei:
        engine.partition_tx_device(0, 0, 0, 0, 0, 1)
    assert ei.value.status == pkg.abi.ERR_INVALID
    with pytest.raises(pkg.IgsError):
        engine.partition_tx_device(0, 0, 257, 0, 0, 1)
