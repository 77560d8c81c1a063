"""World-size-2 CPU test (gloo) of the multi-GPU HOST logic: user-range sharding, the Philox counter base that makes
negatives independent of the rank count, and the two reconciliation formulas (mean of the replicas = the default, sum
of the rank deltas) restated in heat_b200/shard.py.  It pins the formulas and the sharding, not comm.cu: the exchange
kernel itself is tested on GPUs by tests/test_multigpu_gpu.py (bit-identical to the NCCL path and to the single-device
emulation group, replicas identical, 2-rank convergence against one GPU)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from heat_b200 import shard, synth

WORLD = 2


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    d = synth.generate(101, 150, 3000, max_his=20, seed=3, n_clusters=4, max_degree=60)
    D, K = 32, 8
    local, off = shard.shard_clicks(d, rank, WORLD)
    # 1. shards tile the dataset
    sizes = [torch.zeros(3, dtype=torch.int64) for _ in range(WORLD)]
    dist.all_gather(sizes, torch.tensor([local.num_users, local.train_size, off]))
    assert sum(int(s[0]) for s in sizes) == d.num_users and sum(int(s[1]) for s in sizes) == d.train_size
    assert int(sizes[0][2]) == 0 and int(sizes[1][2]) == int(sizes[0][1])
    # 2. negatives of a shard == the same rows of the single-process table
    whole = oracle.neg_table(2022, 1, 0, d.clicks, K, d.num_items)
    mine = oracle.neg_table(2022, 1, off, local.clicks, K, d.num_items)
    ref = whole[off:off + local.train_size]
    # the look-back rule stops at the shard's first sample; only slots whose draw hit the positive there may differ
    assert (mine != ref).mean() < 0.01
    assert np.array_equal(mine[K:], ref[K:]) or (mine != ref).sum() <= K
    # 3. train the shard on the CPU oracle from the shared initial tables, reconcile the item replicas
    uw, iw, w0 = synth.init_tables(d.num_users, d.num_items, D, seed=1)
    u0, u1 = shard.user_range(d.num_users, rank, WORLD)
    my_uw, my_iw, my_w0 = np.ascontiguousarray(uw[u0:u1]), iw.copy(), w0.copy()
    o = oracle.COracle(D, K, local.clicks, local.his_items, local.masks, my_uw, my_iw, my_w0)
    loss = o.train_epoch(seed=2022, sample_offset=off)
    delta = torch.from_numpy(my_iw - iw)
    dist.all_reduce(delta)                                   # what ncclAllReduce(sum) does on the GPU path
    merged = iw + delta.numpy()
    gathered = [torch.zeros_like(torch.from_numpy(my_iw)) for _ in range(WORLD)]
    dist.all_gather(gathered, torch.from_numpy(my_iw))
    want = shard.reconcile(iw, [g.numpy() for g in gathered])
    assert np.allclose(merged, want, rtol=0, atol=1e-7)
    # the default rule: every replica becomes the mean of the replicas == base + mean of the deltas
    mean = torch.from_numpy(my_iw.copy())
    dist.all_reduce(mean)
    assert np.allclose(mean.numpy() / WORLD, shard.reconcile(iw, [g.numpy() for g in gathered], average=True), rtol=0, atol=1e-7)
    tot = torch.tensor([loss * local.train_size, float(local.train_size)], dtype=torch.float64)
    dist.all_reduce(tot)                                     # engine.cpp:380-383
    if rank == 0:
        out.put((float(tot[0] / tot[1]), float(np.abs(merged - iw).max())))
    dist.destroy_process_group()


def test_two_rank_sharding_and_reconciliation():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, port, out)) for r in range(WORLD)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    mean_loss, moved = out.get(timeout=5)
    assert 0 < mean_loss < 4 and moved > 0


def test_user_ranges_cover_everything():
    for U in (1, 7, 100, 52643):
        for world in (1, 2, 3, 8):
            edges = [shard.user_range(U, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == U
            assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))
            sizes = [b - a for a, b in edges]
            assert max(sizes) - min(sizes) <= 1
