"""Tiny driver for ncu: a few epochs of one engine on a BASELINE shape.  profile_train.py <shape> <engine> [epochs] [kw=val ...]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from heat_b200 import abi, synth

name, engine = sys.argv[1], int(sys.argv[2])
epochs = int(sys.argv[3]) if len(sys.argv) > 3 else 2
kw = {k: int(v) for k, v in (a.split("=") for a in sys.argv[4:])}
d = synth.generate_named(name)
t = synth.init_tables(d.num_users, d.num_items, 64)
ctx = abi.Context(64, 16, d.num_users, d.num_items, engine=engine, **kw)
ctx.set_dataset(d.clicks, d.his_items, d.masks)
ctx.bind_tables(*t)
for _ in range(epochs):
    loss = ctx.train_epoch()
    print("loss", loss, "kernel ms", ctx.last_kernel_ms(), flush=True)
