"""Scratch experiment: loss trajectory on the Amazon-Books shape as a function of Hogwild concurrency.

The compiled reference (1 and 8 OpenMP threads) gives 1.038 / 0.608 / 0.533 for epochs 0..2 on this
dataset; this script prints the GPU's trajectory at several grid sizes."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from heat_b200 import abi, synth

name = sys.argv[1] if len(sys.argv) > 1 else "amazon-book"
d = synth.generate_named(name)
D, K = 64, 16
cases = [dict(warps_per_block=12), dict(warps_per_block=12, history_mode=1), dict(warps_per_block=12, update_mode=2),
         dict(warps_per_block=8), dict(warps_per_block=4), dict(warps_per_block=4, max_blocks=32), dict(warps_per_block=4, max_blocks=2)]
for kw in cases:
    t = synth.init_tables(d.num_users, d.num_items, D)
    ctx = abi.Context(D, K, d.num_users, d.num_items, **kw)
    ctx.set_dataset(d.clicks, d.his_items, d.masks)
    ctx.bind_tables(*t)
    out = []
    for ep in range(4):
        out.append((round(ctx.train_epoch(), 4), round(ctx.last_kernel_ms(), 1)))
    print(name, kw, out, flush=True)
    ctx.close()
