"""Speed / recall trade-off of the hot-tile negative sampler (the paper's Table 6 experiment) on a
BASELINE shape: uniform sampler vs tile_sampler with (tile_size, refresh_interval) settings, default engine and the
warp-per-sample engine.  Usage: tile_tradeoff.py [shape] [epochs] > gpurun_out/tile_tradeoff.json"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from heat_b200 import abi, synth

name = sys.argv[1] if len(sys.argv) > 1 else "amazon-book"
epochs = int(sys.argv[2]) if len(sys.argv) > 2 else 5
D, K = 64, 16
d = synth.generate_named(name)
rows = []
for engine in (0, 1):
    for tile in (None, (512, 8192), (4096, 8192), (512, 512)):
        t = synth.init_tables(d.num_users, d.num_items, D)
        kw = dict(neg_sampler=1, tile_sampler=1, tile_size=tile[0], refresh_interval=tile[1]) if tile else {}
        ctx = abi.Context(D, K, d.num_users, d.num_items, engine=engine, **kw)
        ctx.set_dataset(d.clicks, d.his_items, d.masks)
        ctx.bind_tables(*t)
        losses, ms = [], []
        for _ in range(epochs):
            losses.append(ctx.train_epoch()); ms.append(ctx.last_kernel_ms())
        res, _ = ctx.eval_metrics(d.train_indptr, d.train_items, d.test_indptr, d.test_items)
        ctx.close()
        rows.append({"shape": name, "engine": engine, "tile": tile, "kernel_ms": float(np.mean(ms[1:])), "losses": losses,
                     "Recall(k=20)": res["Recall(k=20)"], "NDCG(k=20)": res["NDCG(k=20)"]})
        print(json.dumps(rows[-1]), flush=True)
