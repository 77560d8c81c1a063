"""User-range sharding for one-process-per-GPU training (host-side logic, no GPU needed).

Mirrors what the reference's driver intends (cf_cpu/cf/main.py:51-66 + datasets.py:110-158):
users are split into `world` contiguous id ranges, k = U // world, r = U % world, rank i owns
[i*k + min(i, r), i*k + min(i, r) + k + (i < r)); user ids are re-based to the local range, item
ids and the item count stay global.  (The reference's own rank-0 range, `k + min(0, r)`,
main.py:65, drops the remainder user for rank 0 when r > 0; the formula above is the one its
loop uses for every other rank, applied uniformly.)  Unlike SubClickDataset, the history rows are
sliced from the parent, not re-sampled with an unseeded RNG, so a sharded run sees exactly the
single-process arrays.
"""
from __future__ import annotations

import numpy as np

from .synth import ClickData


def user_range(num_users: int, rank: int, world: int) -> tuple[int, int]:
    k, r = divmod(num_users, world)
    start = rank * k + min(rank, r)
    return start, start + k + (1 if rank < r else 0)


def shard_clicks(d: ClickData, rank: int, world: int):
    """Returns (local ClickData, sample_offset): the rank's users, their samples in file order with
    local user ids, and the global index of its first sample (the Philox counter base, so that
    every sample draws the same negatives whatever the number of ranks)."""
    u0, u1 = user_range(d.num_users, rank, world)
    s0, s1 = int(d.train_indptr[u0]), int(d.train_indptr[u1])
    clicks = d.clicks[s0:s1].copy()
    clicks[:, 0] -= np.uint64(u0)
    t0, t1 = int(d.test_indptr[u0]), int(d.test_indptr[u1])
    local = ClickData(
        num_users=u1 - u0, num_items=d.num_items, max_his=d.max_his, clicks=np.ascontiguousarray(clicks),
        his_items=np.ascontiguousarray(d.his_items[u0:u1]), masks=np.ascontiguousarray(d.masks[u0:u1]),
        train_indptr=(d.train_indptr[u0:u1 + 1] - s0).astype(np.int64), train_items=d.train_items[s0:s1].copy(),
        test_indptr=(d.test_indptr[u0:u1 + 1] - t0).astype(np.int64), test_items=d.test_items[t0:t1].copy())
    return local, s0


def reconcile(base: np.ndarray, replicas: list[np.ndarray], average: bool = False) -> np.ndarray:
    """Host restatement of heat_comm_sync_items: base + sum (or mean) of the replicas' deltas."""
    delta = np.zeros_like(base, dtype=np.float32)
    for w in replicas:
        delta += (w - base).astype(np.float32)
    if average:
        delta *= np.float32(1.0 / len(replicas))
    return (base + delta).astype(np.float32)
