"""Which shared state makes the loss / Recall depend on the number of samples in flight?

    [HEAT_EXP_LR_W0_SCALE=0] python tools/concurrency_exp.py <shape> <epochs> case [case ...]

A case is `engine:warps:max_blocks:update_mode[:chunk]` (engine 1 = warp per sample, 2 = CTA per sample; 0 = default
for a field).  Prints per-epoch losses, kernel ms and Recall@20 of every case on the same data and initial tables, so
that two runs of the script -- with the aggregator's W0 learning (HEAT_EXP_LR_W0_SCALE unset) and with W0 frozen
(=0) -- separate the contribution of the dense shared W0 from that of the item rows."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from heat_b200 import abi, synth

name, epochs = sys.argv[1], int(sys.argv[2])
d = synth.generate_named(name)
D, K = 64, 16
clip = 0.1 if name == "gowalla" else 1.0
print("HEAT_EXP_LR_W0_SCALE =", os.environ.get("HEAT_EXP_LR_W0_SCALE", "(unset: 1)"), flush=True)
for case in sys.argv[3:]:
    f = [int(x) for x in case.split(":")] + [0] * 5
    t = synth.init_tables(d.num_users, d.num_items, D)
    ctx = abi.Context(D, K, d.num_users, d.num_items, clip_val=clip, engine=f[0], warps_per_block=f[1], max_blocks=f[2],
                      update_mode=f[3], chunk=f[4])
    ctx.set_dataset(d.clicks, d.his_items, d.masks)
    ctx.bind_tables(*t)
    losses, ms = [], []
    for ep in range(epochs):
        losses.append(round(ctx.train_epoch(), 4)); ms.append(ctx.last_kernel_ms())
    res, _ = ctx.eval_metrics(d.train_indptr, d.train_items, d.test_indptr, d.test_items)
    print(f"{name} {case:16s} loss {losses} R@20 {res['Recall(k=20)']:.4f} {np.mean(ms):.1f} ms/epoch", flush=True)
    ctx.close()
