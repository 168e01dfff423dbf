"""Linked accounts across shards: drives the three device-side stages of
`igs_links_*` (include/igs.h) and the two all-to-all exchanges between them.

This is the only place of the path where data moves between GPUs: accounts are sharded by account,
devices are not, so the (device, account) pairs travel to the shard that owns the device, and the links
it finds travel to the shards that own the accounts.  One process per GPU; the exchange is
`torch.distributed.all_to_all_single` (NCCL over NVLink / NVSwitch on device buffers; gloo on host
copies in the CPU tests).  With one shard nothing is exchanged and torch is not imported.

Plumbing only: every record is produced and consumed by the CUDA library.
"""
from __future__ import annotations

from typing import Optional

import numpy as np

RECORD_BYTES = 16        # igs_link_pair and igs_link


class _DeviceBytes:
    """A device buffer owned by the engine, exposed to torch through __cuda_array_interface__."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def _exchange(d_ptr: int, counts: np.ndarray, device: int, group):
    """all-to-all of grouped 16-byte records: rank r sends group s of its buffer to rank s.
    -> (tensor holding the received records (kept alive by the caller), its device pointer, record count)"""
    import torch
    import torch.distributed as dist

    dev = torch.device("cuda", device)
    n_shards = counts.shape[0]
    send_counts = torch.from_numpy(counts.astype(np.int64)).to(dev)
    recv_counts = torch.empty(n_shards, dtype=torch.int64, device=dev)
    dist.all_to_all_single(recv_counts, send_counts, group=group)
    recv = recv_counts.cpu().numpy()
    total_send = int(counts.sum())
    send = (torch.as_tensor(_DeviceBytes(d_ptr, total_send * RECORD_BYTES), device=dev) if total_send
            else torch.empty(0, dtype=torch.uint8, device=dev))
    out = torch.empty(int(recv.sum()) * RECORD_BYTES, dtype=torch.uint8, device=dev)
    dist.all_to_all_single(out, send, output_split_sizes=[int(c) * RECORD_BYTES for c in recv],
                           input_split_sizes=[int(c) * RECORD_BYTES for c in counts], group=group)
    torch.cuda.synchronize(dev)
    return out, (out.data_ptr() if out.numel() else 0), int(recv.sum())


def build_links(engine, shard: int = 0, n_shards: int = 1, group=None, exchange=None) -> dict:
    """Rebuilds the per-account links table of `engine`'s shard from the loaded history (a periodic job,
    like the hourly batch features).  Collective: every shard's process calls it.  `exchange` overrides the
    all-to-all (tests); it gets (device pointer, per-shard counts) and returns (keep-alive, pointer, count).
    Returns the record counts of each stage."""
    if exchange is None:
        if n_shards == 1:
            exchange = lambda ptr, counts: (None, ptr, int(counts.sum()))       # noqa: E731
        else:
            exchange = lambda ptr, counts: _exchange(ptr, counts, engine.device, group)   # noqa: E731
    d_pairs, pair_counts = engine.links_pairs(shard, n_shards)
    keep1, d_recv_pairs, n_pairs = exchange(d_pairs, pair_counts)
    d_links, link_counts = engine.links_group(d_recv_pairs, n_pairs, n_shards)
    keep2, d_recv_links, n_links = exchange(d_links, link_counts)
    engine.links_finish(d_recv_links, n_links, shard)
    del keep1, keep2
    return {"pairs_sent": int(pair_counts.sum()), "pairs_received": n_pairs,
            "links_sent": int(link_counts.sum()), "links_received": n_links}


def setup_peer_windows(engine, shard: int, n_shards: int, cap_pairs: int, cap_links: int, group=None) -> None:
    """Collective, once: every shard allocates its receive window and maps every peer's (CUDA IPC handles
    travel through torch.distributed; NVLink / NVSwitch carries the stores and atomics afterwards)."""
    handle, ptr = engine.links_window_create(cap_pairs, cap_links)
    if n_shards > 1:
        import torch.distributed as dist
        handles = [None] * n_shards
        dist.all_gather_object(handles, handle, group=group)
        for s in range(n_shards):
            if s != shard:
                engine.links_window_open(s, handles[s])
    engine.links_window_attach(shard, ptr)


def build_links_peer(engine, shard: int = 0, n_shards: int = 1, group=None, barrier=None) -> None:
    """build_links with both exchanges fused into the producing kernels over peer memory (no NCCL on the
    data, no staging, no host-side counts).  Needs setup_peer_windows.  No kernel waits for another GPU: the
    steps are separated by host barriers."""
    if barrier is None:
        if n_shards == 1:
            barrier = lambda: None                                           # noqa: E731
        else:
            import torch.distributed as dist
            barrier = lambda: dist.barrier(group=group)                      # noqa: E731
    engine.links_window_reset()
    barrier()
    engine.links_pairs_peer(shard, n_shards)
    barrier()
    engine.links_group_peer(shard, n_shards)
    barrier()
    engine.links_finish_peer(shard)


def global_account_id(shard: int, index) -> np.ndarray:
    """(shard << 32) | dense index — the account ids the links table reports."""
    return (np.uint64(shard) << np.uint64(32)) | np.asarray(index, dtype=np.uint64)
