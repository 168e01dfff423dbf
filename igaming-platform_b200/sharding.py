"""Host-side shard bookkeeping for N GPUs: the path shards by account, every rank scores
its shard independently and only results are gathered — there is no collective on the data path.

`shard_of` is the numpy twin of the device rule in csrc/k_partition.cu and of igs_shard_of(a, 0, n)."""
from __future__ import annotations

from typing import List, Tuple

import numpy as np


def splitmix64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        x = x.astype(np.uint64) + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return x ^ (x >> np.uint64(31))


def shard_of(account: np.ndarray, n_shards: int) -> np.ndarray:
    return (splitmix64(account) % np.uint64(n_shards)).astype(np.uint32)


def split_by_shard(records: np.ndarray, n_shards: int, key: str = "account") -> Tuple[List[np.ndarray], np.ndarray, np.ndarray]:
    """Stable grouping of `records` by shard: (per-shard record arrays, perm, offsets) with
    grouped = records[perm], shard s = grouped[offsets[s]:offsets[s+1]] — what the device partition
    (igs_partition_tx_device) returns."""
    sh = shard_of(records[key], n_shards)
    perm = np.argsort(sh, kind="stable").astype(np.uint32)
    offsets = np.concatenate([[0], np.cumsum(np.bincount(sh, minlength=n_shards))]).astype(np.uint32)
    grouped = records[perm]
    return [grouped[offsets[s]:offsets[s + 1]] for s in range(n_shards)], perm, offsets


def merge_results(per_shard: List[np.ndarray], perm: np.ndarray) -> np.ndarray:
    """Inverse of split_by_shard for the per-shard result arrays: results in the original order."""
    grouped = np.concatenate(per_shard) if per_shard else np.zeros(0)
    out = np.empty_like(grouped)
    out[perm] = grouped
    return out


def local_index(accounts: np.ndarray, n_shards: int) -> Tuple[np.ndarray, np.ndarray]:
    """Dense per-shard index of global account ids 0..A-1: (shard, local index) of every account, with
    local indices assigned in increasing global id inside a shard."""
    sh = shard_of(accounts, n_shards)
    local = np.zeros(accounts.shape[0], dtype=np.uint32)
    for s in range(n_shards):
        m = sh == s
        local[m] = np.arange(int(m.sum()), dtype=np.uint32)
    return sh, local
