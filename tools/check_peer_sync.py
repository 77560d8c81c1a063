"""Multi-GPU check of the fused peer-memory exchange (comm.cu: peer_sync_kernel): against the NCCL path, and against the
single-device rank emulation (heat_local_group_*).  Run by tests/test_multigpu_gpu.py when >= 2 GPUs are visible.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 tools/check_peer_sync.py

Every rank trains its own small shard twice from identical tables in deterministic mode (one block, file order:
bit-reproducible), once with exchange_mode 0 (fused kernel over CUDA-IPC peer memory) and once with exchange_mode 1
(NCCL all-reduce path), two exchanges per epoch, two epochs, for both reconciliation rules (mean of the replicas /
sum of the rank deltas).  At N = 2 a sum has two addends, so the two paths must agree BIT FOR BIT; at N > 2 NCCL's
summation order differs and they must agree to 1e-6 relative.  The fused kernel adds in rank order, exactly as the
emulation group does, so rank 0 re-runs all N ranks as an emulation group on its own GPU and the tables must be
bit-identical at every N.  All ranks must end with identical item tables.  The exchange time of both paths at the
Amazon-Books table size is printed."""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from heat_b200 import abi, synth

N_SAMPLES = 12000


def shard_of(q):
    return synth.generate(600, 900, N_SAMPLES, seed=100 + q)          # rank q: its own users, the shared item ids


def make(rank, world, mode, d, tables, connect=True, device=None, **kw):
    ctx = abi.Context(64, 16, d.num_users, d.num_items, seed=2022, sample_offset=rank * d.train_size,
                      device=rank if device is None else device, exchange_mode=mode, **kw)
    ctx.set_dataset(d.clicks, d.his_items, d.masks)
    ctx.bind_tables(*tables)
    if connect:
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.from_numpy(abi.Context.comm_unique_id()))
        dist.broadcast(uid, 0)
        ctx.comm_init(rank, world, uid.cpu().numpy())
    return ctx


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
    d = shard_of(rank)
    init = synth.init_tables(d.num_users, d.num_items, 64, seed=7)
    interval = N_SAMPLES // 2 + 1
    for avg, grads in ((1, 0), (2, 0), (2, 1)):
        out = {}
        for mode in (0, 1):
            t = [a.copy() for a in init]
            ctx = make(rank, world, mode, d, t, deterministic=True, sync_average=avg, sync_grads=grads)
            ctx.set_sync_interval(interval)
            want = 2 if mode == 0 else 1
            assert ctx.comm_mode() == want, f"rank {rank}: comm_mode {ctx.comm_mode()} (expected {want})"
            losses = [ctx.train_epoch() for _ in range(2)]
            assert ctx.last_exchange()[0] == 2
            ctx.sync_to_host()
            out[mode] = (t[1].copy(), t[2].copy(), losses)
            ctx.close()
        iw0, w00, l0 = out[0]
        iw1, w01, l1 = out[1]
        if world == 2:
            assert np.array_equal(iw0, iw1) and np.array_equal(w00, w01), "fused exchange differs from the NCCL path"
            assert l0 == l1
        else:
            assert np.abs(iw0 - iw1).max() <= 1e-6 * np.abs(iw1).max() and np.abs(w00 - w01).max() <= 1e-6 * np.abs(w01).max()
        g = torch.from_numpy(iw0).cuda()
        ref = g.clone()
        dist.broadcast(ref, 0)
        assert torch.equal(g, ref), "replicas differ after the fused exchange"
        if rank == 0:           # the same N ranks as an emulation group on this GPU: same arithmetic, same order
            ctxs, keep = [], []
            for q in range(world):
                dq = shard_of(q)
                tq = [a.copy() for a in synth.init_tables(dq.num_users, dq.num_items, 64, seed=7)]
                keep.append(tq)
                c = make(q, world, 0, dq, tq, connect=False, device=0, deterministic=True, sync_average=avg, sync_grads=grads)
                c.set_sync_interval(interval)
                ctxs.append(c)
            grp = abi.LocalGroup(ctxs)
            el = [grp.train_epoch() for _ in range(2)]
            ctxs[0].sync_to_host()
            assert np.array_equal(keep[0][1], iw0) and np.array_equal(keep[0][2], w00), "emulation group differs from the real ranks"
            assert np.allclose(el, l0, rtol=1e-12), (el, l0)
            for c in ctxs:
                c.close()
        dist.barrier()

    # exchange time by table size (91 599 x 64 fp32 = the Amazon-Books table; 2 M x 64 = BASELINE configs[4]), both paths,
    # default rule (mean).  Back-to-back exchanges, wall clock over 20 of them; NVLink bytes per rank and exchange =
    # (world-1)/world of the table read from the peers + the same written to them.
    for items in (91599, 500000, 2000000):
        big = synth.generate(64, items, 4096, seed=3, max_degree=64)
        for mode in (0, 1):
            t = list(synth.init_tables(big.num_users, big.num_items, 64, seed=7))
            ctx = make(rank, world, mode, big, t)
            for _ in range(3):
                ctx.comm_sync_items()
            torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(20):
                ctx.comm_sync_items()
            ctx.sync_to_host()
            dist.barrier()
            ms = (time.perf_counter() - t0) / 20 * 1e3
            mb = items * 64 * 4 / 1e6
            if rank == 0:
                link = 2 * (world - 1) / world * mb / 1e3 / (ms * 1e-3)
                print(f"exchange of {mb:.1f} MB at world {world}: {'fused peer kernel' if mode == 0 else 'NCCL all-reduce path'} {ms:.3f} ms"
                      f" = {link:.0f} GB/s of NVLink traffic per rank (read + write)", flush=True)
            ctx.close()
    if rank == 0:
        print(f"peer exchange check ok at world {world}", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
