"""Per-epoch loss and final Recall/NDCG of the GPU path vs the compiled reference (oracle/_ref) on the
same seeded synthetic inputs (BASELINE configs 1-3).  Usage: recall_parity.py <shape> <epochs> [modes...]
A mode is `auto` (library defaults), `cta:<max_blocks>:<blocks_per_sm>` (engine 2) or `warps:max_blocks:update_mode` (engine 1;
update_mode 0/1 fresh-grad, 2 atomic, 3 snapshot store), e.g. 12:0:0 (full grid) or 4:32:3 (128 warps in flight, store)."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from heat_b200 import abi, synth

name, epochs = sys.argv[1], int(sys.argv[2])
modes = sys.argv[3:] or ["12:0:0", "4:0:0", "4:32:0"]
D, K = 64, 16
clip = 0.1 if name == "gowalla" else 1.0          # the YAMLs' clip_val (Gowalla 0.1, others 1.0)
d = synth.generate_named(name)
out = {"shape": name, "epochs": epochs, "clip": clip, "runs": []}


def evaluate(tables):
    ctx = abi.Context(D, K, d.num_users, d.num_items)
    ctx.bind_tables(*tables)
    res, _ = ctx.eval_metrics(d.train_indptr, d.train_items, d.test_indptr, d.test_items)
    ms = ctx.last_eval_ms()
    ctx.close()
    return res, ms


t = synth.init_tables(d.num_users, d.num_items, D)
o = oracle.RefOracle(D, K, d.clicks, d.his_items, d.masks, *t, clip_val=clip)
t0 = time.time()
ref_losses = [o.train_epoch(nthreads=0, sampler=2) for _ in range(epochs)]      # same Philox negatives as the GPU
ref_s = (time.time() - t0) / epochs
ref_res, _ = evaluate(t)
out["reference"] = {"losses": ref_losses, "sec_per_epoch": ref_s, "threads": o.max_threads(), **ref_res}
print("reference", [round(x, 4) for x in ref_losses], {k: round(v, 4) for k, v in ref_res.items()}, f"{ref_s:.2f} s/epoch", flush=True)
for mode in modes:
    t = synth.init_tables(d.num_users, d.num_items, D)
    if mode == "auto":                    # library defaults
        ctx = abi.Context(D, K, d.num_users, d.num_items, clip_val=clip)
    elif mode == "strict":                # engine 2 at its default concurrency: blocks = min(#SMs, num_items/600)
        ctx = abi.Context(D, K, d.num_users, d.num_items, clip_val=clip, engine=2)
    elif mode.startswith("cta"):          # cta:<max_blocks>:<blocks_per_sm>
        _, mb, bps = mode.split(":")
        ctx = abi.Context(D, K, d.num_users, d.num_items, clip_val=clip, engine=2, max_blocks=int(mb), blocks_per_sm=int(bps))
    else:                                 # <warps>:<max_blocks>:<update_mode>  (warp-per-sample engine)
        w, mb, upd = map(int, mode.split(":"))
        ctx = abi.Context(D, K, d.num_users, d.num_items, clip_val=clip, engine=1, warps_per_block=w, max_blocks=mb, update_mode=upd)
    ctx.set_dataset(d.clicks, d.his_items, d.masks)
    ctx.bind_tables(*t)
    losses, ms = [], []
    for _ in range(epochs):
        losses.append(ctx.train_epoch()); ms.append(ctx.last_kernel_ms())
    ctx.sync_to_host()
    ctx.close()
    res, ems = evaluate(t)
    run = {"mode": mode, "losses": losses, "kernel_ms": float(np.mean(ms)), "eval_ms": ems, **res,
           "dRecall20": res["Recall(k=20)"] - ref_res["Recall(k=20)"], "dLossLast": losses[-1] - ref_losses[-1]}
    out["runs"].append(run)
    print(mode, [round(x, 4) for x in losses], {k: round(v, 4) for k, v in res.items()},
          f"{run['kernel_ms']:.1f} ms/epoch, eval {ems:.1f} ms, dRecall@20 {run['dRecall20']:+.4f}", flush=True)
os.makedirs("gpurun_out", exist_ok=True)
with open(f"gpurun_out/recall_parity_{name}_{epochs}.json", "w") as f:
    json.dump(out, f, indent=1)
