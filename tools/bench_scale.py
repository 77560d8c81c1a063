"""BASELINE.json configs[4]: 10 M users x 2 M items x 500 M interactions on 8 B200 -- user-sharded training + row-sharded
evaluation (score GEMM / top-k with the seven metric sums all-reduced).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 \
        tools/bench_scale.py [--scale 1] [--epochs 3] > gpurun_out/cfg5_n8.json
    python tools/bench_scale.py            # one rank = ONE GPU's shard of the same dataset (block 0)

Rank r generates and trains user block r of ONE structured power-law dataset (heat_b200.synth.generate_powerlaw_block:
shared item clusters and popularity law, 1.25 M users and ~60 M samples per block at --scale 1), against its replica of
the 2 M-item table; the replicas are reconciled by the library's default exchange (mean of the replicas every
max(65536, 8 x items) samples, one kernel per rank over NVLink peer memory).  One JSON line (rank 0): per-rank and
total samples/s, exchanges and their share of the epoch, evaluation ms, Recall, and the size-independent parity
properties (sampler slice bit-exact with the host twin, finite decreasing loss, top-k rows equal to an fp64 host
recomputation on sampled users).  --scale S divides users, items and samples by S; with --reference (small scales) rank 0
also trains the UNION of all blocks with the CPU reference (oracle/_ref) and reports the Recall difference."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import bench
    from heat_b200 import abi, synth

    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=int, default=1)
    ap.add_argument("--epochs", type=int, default=3)
    ap.add_argument("--engine", type=int, default=0)
    ap.add_argument("--max-blocks", type=int, default=0)
    ap.add_argument("--sync-interval", type=int, default=0)
    ap.add_argument("--no-eval", action="store_true")
    ap.add_argument("--reference", action="store_true", help="rank 0 trains the union with the CPU reference too (small --scale only)")
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    D, K = 64, 16
    users, items, samples = 1_250_000 // args.scale, 2_000_000 // args.scale, 62_500_000 // args.scale
    t0 = time.perf_counter()
    d = synth.generate_powerlaw_block(rank, users, items, samples, seed=2022)
    def user_rows(q):                                   # block q's initial user rows: (seed, q) alone, like its clicks
        return (np.random.default_rng([2022, 104729, q]).standard_normal((users, D)) * 1e-2).astype(np.float32)
    _, iw, w0 = synth.init_tables(1, items, D, seed=2022)            # item table / W0: identical on every rank
    uw = user_rows(rank)
    gen_s = time.perf_counter() - t0
    N = d.train_size
    offs = np.zeros(world + 1, np.int64)
    if world > 1:
        t = torch.zeros(world, dtype=torch.int64, device="cuda"); t[rank] = N
        dist.all_reduce(t)
        offs[1:] = np.cumsum(t.cpu().numpy())
    else:
        offs[1] = N
    ctx = abi.Context(D, K, d.num_users, d.num_items, clip_val=1.0, l_r=1e-2, milestones=(10,), seed=2022, sample_offset=int(offs[rank]),
                      engine=args.engine, max_blocks=args.max_blocks, device=local)
    t0 = time.perf_counter()
    ctx.set_dataset(d.clicks, d.his_items, d.masks)
    ctx.bind_tables(uw, iw, w0)
    upload_s = time.perf_counter() - t0
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.from_numpy(abi.Context.comm_unique_id()))
        dist.broadcast(uid, 0)
        ctx.comm_init(rank, world, uid.cpu().numpy())
        ctx.set_sync_interval(args.sync_interval)
    alg = bench.algorithmic_bytes(d.clicks, d.masks, D, K)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def rmax(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    ms, kms, xms, losses, nx = [], [], [], [], 0
    for _ in range(args.epochs):
        flush.add_(1); torch.cuda.synchronize()
        if world > 1:
            dist.barrier(); torch.cuda.synchronize()
        losses.append(ctx.train_epoch())
        ms.append(rmax(ctx.last_epoch_ms())); kms.append(ctx.last_kernel_ms())
        nx, x = ctx.last_exchange(); xms.append(x)
    s0 = N // 2
    negs = ctx.sample_negatives(1, s0, s0 + 256)
    pos = d.clicks[s0:s0 + 256, 1]
    twin = np.array([[abi.philox_draw(2022, 1, int(offs[rank]) + s0 + i, k, d.num_items) for k in range(K)] for i in range(256)], dtype=np.uint64)
    clash = twin == pos[:, None]                                    # "draw == positive" slots take the look-back path
    total_n = int(offs[-1])
    steady = ms[1:] or ms
    row = {"config": f"BASELINE.json configs[4] / scale {args.scale}: {world} rank(s), user blocks of one structured power-law dataset",
           "world": world, "users_per_rank": d.num_users, "items": d.num_items, "samples_rank0": N, "samples_total": total_n,
           "emb_dim": D, "num_negs": K, "table_bytes_w_plus_g_per_rank": 2 * (d.num_users + d.num_items) * D * 4,
           "host_generate_s": gen_s, "upload_s": upload_s, "epoch_ms_max_over_ranks": ms, "kernel_ms_rank0": kms,
           "samples_per_s_total": total_n / (float(np.mean(steady)) * 1e-3), "samples_per_s_per_rank": total_n / world / (float(np.mean(steady)) * 1e-3),
           "algorithmic_bytes_per_epoch_rank0": alg, "algorithmic_GBps_rank0": alg / (float(np.mean(kms[1:] or kms)) * 1e-3) / 1e9,
           "exchanges_per_epoch": int(nx), "exchange_ms_per_epoch_rank0": float(np.mean(xms[1:] or xms)),
           "exchange_share_of_epoch": float(np.mean(xms[1:] or xms)) / float(np.mean(steady)), "exchange": ctx.comm_mode(),
           "exchanged_bytes_per_exchange": (d.num_items * D + D * D) * 4,
           "losses": losses, "loss_decreasing": bool(all(b < a for a, b in zip(losses, losses[1:]))), "loss_finite": bool(np.isfinite(losses).all()),
           "sampler_slice_bit_exact": bool(np.array_equal(negs[~clash], twin[~clash])), "engine": ctx.last_engine()}
    if not args.no_eval:
        t0 = time.perf_counter()
        res, sums = ctx.eval_metrics(d.train_indptr, d.train_items, d.test_indptr, d.test_items)
        wall = (time.perf_counter() - t0) * 1e3
        ev = rmax(ctx.last_eval_ms())
        sums = ctx.allreduce_f64(sums)
        row["eval_metrics_ms_max_over_ranks"] = ev
        row["eval_wall_ms_rank0"] = wall
        row["eval_tf32_tflops_executed_per_rank"] = 3 * 2.0 * d.num_users * d.num_items * D / (ctx.last_eval_ms() * 1e-3) / 1e12
        row["metrics_all_users"] = dict(zip(res.keys(), (sums[:6] / max(sums[6], 1.0)).tolist()))
        ids, sc = ctx.eval_topk(50, d.train_indptr, d.train_items)
        ctx.sync_to_host()
        rng = np.random.default_rng(0)
        ok = True
        iw64 = iw.astype(np.float64)
        for u in rng.integers(0, d.num_users, 16):
            s = iw64 @ uw[u].astype(np.float64)
            tr = d.train_items[d.train_indptr[u]:d.train_indptr[u + 1]]
            s[tr] = -np.inf
            want = np.lexsort((np.arange(d.num_items), -s))[:50]
            got = ids[u]
            same = set(want.tolist()) == set(got.tolist()) or np.allclose(np.sort(s[got])[::-1], s[want], rtol=1e-5, atol=1e-7)
            ok &= bool(same) and bool(np.all(np.diff(sc[u]) <= 0)) and not np.isin(got, tr).any()
        row["topk_rows_match_host_fp64"] = ok
    ctx.close()
    if args.reference and rank == 0:
        import oracle
        blocks = [d] + [synth.generate_powerlaw_block(q, users, items, samples, seed=2022) for q in range(1, world)]
        un = synth.concat_blocks(blocks)
        _, i0, w00 = synth.init_tables(1, items, D, seed=2022)
        t = (np.concatenate([user_rows(q) for q in range(world)]), i0, w00)
        o = oracle.RefOracle(D, K, un.clicks, un.his_items, un.masks, *t) if oracle.RefOracle.available() else \
            oracle.COracle(D, K, un.clicks, un.his_items, un.masks, *t)
        thr = max(1, len(os.sched_getaffinity(0)))
        rl = [o.train_epoch(nthreads=thr, sampler=2) if oracle.RefOracle.available() else o.train_epoch(nthreads=thr) for _ in range(args.epochs)]
        o.close()
        c = abi.Context(D, K, un.num_users, un.num_items, device=local)
        c.bind_tables(*t)
        rr, _ = c.eval_metrics(un.train_indptr, un.train_items, un.test_indptr, un.test_items)
        c.close()
        row["reference_union"] = {"losses": rl, "threads": thr, **rr}
        if "metrics_all_users" in row:
            row["recall20_vs_reference"] = row["metrics_all_users"]["Recall(k=20)"] - rr["Recall(k=20)"]
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
