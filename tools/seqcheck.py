"""Deterministic GPU run of the first n samples vs the oracle stepping the same samples with the same negatives:
growth of the difference with n (smooth = fp32 rounding amplified by the dynamics; a jump = an ordering bug)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, oracle
from heat_b200 import abi, synth
from helpers import rel_err
name = sys.argv[1] if len(sys.argv) > 1 else "small"
engine = int(sys.argv[2]) if len(sys.argv) > 2 else 0
d = synth.generate_named(name)
D, K = 64, 16
ns = [1, 3, 10, 30, 33, 64, 70, 100, 130, 200, 300, 600, 1000]
a = synth.init_tables(d.num_users, d.num_items, D)
ctx = abi.Context(D, K, d.num_users, d.num_items, deterministic=True, w0_tail_mode=1, engine=engine)
ctx.set_dataset(d.clicks, d.his_items, d.masks); ctx.bind_tables(*a)
negs = ctx.sample_negatives(0, 0, max(ns))
b = synth.init_tables(d.num_users, d.num_items, D)
o = oracle.COracle(D, K, d.clicks, d.his_items, d.masks, *b)
done = 0
for n in ns:
    init = synth.init_tables(d.num_users, d.num_items, D)
    for t, t0 in zip(a, init):
        t[...] = t0
    ctx.upload_tables(); ctx.zero_grad()
    _, per = ctx.train_range(0, n, per_sample=True)
    ctx.sync_to_host()
    ref = []
    while done < n:
        ref.append(o.step(int(d.clicks[done, 0]), int(d.clicks[done, 1]), negs[done])); done += 1
    print(name, "n", n, "user", f"{rel_err(a[0], b[0]):.2e}", "item", f"{rel_err(a[1], b[1]):.2e}", "w0", f"{rel_err(a[2], b[2]):.2e}",
          "last loss gpu/oracle", float(per[-1]), ref[-1] if ref else None, flush=True)
