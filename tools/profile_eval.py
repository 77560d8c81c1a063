"""Tiny driver for ncu: evaluation (split + score GEMM/top-k + metrics) on a BASELINE shape.  profile_eval.py <shape> [reps] [train epochs]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from heat_b200 import abi, synth

name = sys.argv[1]
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
epochs = int(sys.argv[3]) if len(sys.argv) > 3 else 1
d = synth.generate_named(name)
t = synth.init_tables(d.num_users, d.num_items, 64)
ctx = abi.Context(64, 16, d.num_users, d.num_items)
ctx.set_dataset(d.clicks, d.his_items, d.masks)
ctx.bind_tables(*t)
for _ in range(epochs):
    ctx.train_epoch()
for _ in range(reps):
    res, _ = ctx.eval_metrics(d.train_indptr, d.train_items, d.test_indptr, d.test_items)
    print({k: round(v, 4) for k, v in res.items()}, "eval ms", ctx.last_eval_ms(), flush=True)
