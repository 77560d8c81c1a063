"""Convergence of user-sharded training against ONE GPU training the union of the shards, on a single device.

    python tools/multigpu_sweep.py <shape> <world> <epochs> [--ref] policy [policy ...]

The dataset is the weak-scaling form of <shape>: <world> x the users over the same item catalogue, one seed.  It is
trained (a) by one context over all users ("n1": what --gpus 1 would do with this data), (b) by <world> contexts, one
user-range shard each, as an emulation group (heat_local_group_*: the ranks' segments one after the other, then the
same reconciliation a real run does).  A policy is `interval:average:grads[:engine[:share]]` -- samples per rank between
exchanges (0 = once per epoch), 0/1 mean of the replicas, 2 sum of the rank deltas, 1 = grad rows exchanged too.  --ref adds the CPU
reference (oracle/_ref, all host threads) on the union.  Prints per-epoch losses and Recall@20 of every arm.
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from heat_b200 import abi, shard, synth

args = [a for a in sys.argv[1:] if not a.startswith("--")]
flags = [a for a in sys.argv[1:] if a.startswith("--")]
name, world, epochs = args[0], int(args[1]), int(args[2])
policies = args[3:]
repeat = max([int(f.split("=")[1]) for f in flags if f.startswith("--repeat=")] or [1])
policies = [p for p in policies for _ in range(repeat)]
D, K = 64, 16
clip = 0.1 if name == "gowalla" else 1.0
t0 = time.time()
d = synth.generate_named(name, users_mult=world)
print(f"union dataset: {d.num_users} users, {d.num_items} items, {d.train_size} samples ({time.time() - t0:.1f} s)", flush=True)
out = {"shape": name, "world": world, "epochs": epochs, "runs": []}


def recall_of(ctxs, shards):
    sums = np.zeros(7)
    for c, sh in zip(ctxs, shards):
        _, s = c.eval_metrics(sh.train_indptr, sh.train_items, sh.test_indptr, sh.test_items)
        sums += s
    return {"Recall@20": sums[0] / sums[6], "Recall@50": sums[1] / sums[6], "NDCG@20": sums[2] / sums[6]}


def report(tag, losses, res, ms, extra=""):
    print(f"{tag:28s} loss {[round(x, 4) for x in losses]} R@20 {res['Recall@20']:.4f} R@50 {res['Recall@50']:.4f} "
          f"{ms:.1f} ms/epoch {extra}", flush=True)


# (a) one context over the union
t = synth.init_tables(d.num_users, d.num_items, D)
ctx = abi.Context(D, K, d.num_users, d.num_items, clip_val=clip)
ctx.set_dataset(d.clicks, d.his_items, d.masks)
ctx.bind_tables(*t)
losses, ms = [], []
for _ in range(epochs):
    losses.append(ctx.train_epoch()); ms.append(ctx.last_epoch_ms())
n1 = recall_of([ctx], [d])
ctx.close()
report("n1 (one GPU, union)", losses, n1, float(np.mean(ms)))
out["n1"] = {"losses": losses, **n1}

if "--ref" in flags:
    import oracle
    t = synth.init_tables(d.num_users, d.num_items, D)
    o = oracle.RefOracle(D, K, d.clicks, d.his_items, d.masks, *t, clip_val=clip)
    t0 = time.time()
    rl = [o.train_epoch(nthreads=os.cpu_count(), sampler=2) for _ in range(epochs)]
    rs = (time.time() - t0) / epochs
    c = abi.Context(D, K, d.num_users, d.num_items)
    c.bind_tables(*t)
    rr = recall_of([c], [d])
    c.close(); o.close()
    report(f"reference ({os.cpu_count()} threads)", rl, rr, rs * 1e3)
    out["reference"] = {"losses": rl, **rr}

shards = [shard.shard_clicks(d, r, world) for r in range(world)]
for pol in policies:
    f = pol.split(":")
    interval, average, grads = int(f[0]), int(f[1]), int(f[2])
    engine = int(f[3]) if len(f) > 3 else 0
    share = int(f[4]) if len(f) > 4 else 0
    t = synth.init_tables(d.num_users, d.num_items, D)
    ctxs, keep = [], []
    for r, (sh, off) in enumerate(shards):
        u0, u1 = shard.user_range(d.num_users, r, world)
        tabs = (np.ascontiguousarray(t[0][u0:u1]), t[1].copy(), t[2].copy())
        keep.append(tabs)
        c = abi.Context(D, K, sh.num_users, d.num_items, clip_val=clip, sample_offset=off, sync_average=average,
                        sync_grads=grads, engine=engine)
        c.set_dataset(sh.clicks, sh.his_items, sh.masks)
        c.bind_tables(*tabs)
        c.set_sync_interval(interval)
        ctxs.append(c)
    grp = abi.LocalGroup(ctxs, share=share)
    losses = []
    for _ in range(epochs):
        losses.append(grp.train_epoch())
    nx, xms = ctxs[0].last_exchange()
    res = recall_of(ctxs, [s for s, _ in shards])
    kms = ctxs[0].last_kernel_ms()
    for c in ctxs:
        c.close()
    report(f"x{world} {pol}", losses, res, kms, f"| {nx} exchanges {xms:.2f} ms | dR@20 vs n1 {res['Recall@20'] - n1['Recall@20']:+.4f}")
    out["runs"].append({"policy": pol, "losses": losses, "exchanges": nx, "exchange_ms": xms, **res,
                        "dRecall20_vs_n1": res["Recall@20"] - n1["Recall@20"]})
os.makedirs("gpurun_out", exist_ok=True)
with open(f"gpurun_out/multigpu_sweep_{name}_x{world}.json", "w") as fjs:
    json.dump(out, fjs, indent=1)
