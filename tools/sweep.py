"""Quick kernel timing sweep on the Amazon-Books shape (scratch tool, not the bench)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from heat_b200 import abi, synth

d = synth.generate_named("amazon-book")
D, K = 64, 16
H = d.masks[d.clicks[:, 0].astype(np.int64), 0].astype(np.float64)
bytes_epoch = float(np.sum(4 * D * (4 * (2 + K) + H) + 12 + 4 * H))
print("samples", d.train_size, "mean H per sample", H.mean(), "algorithmic GB/epoch", bytes_epoch / 1e9)
print("L2 probe GB/s (74MB):", abi.probe_l2_bandwidth(0, 74 << 20), " (32MB):", abi.probe_l2_bandwidth(0, 32 << 20),
      " (1GB, HBM):", abi.probe_l2_bandwidth(0, 1 << 30))
for warps in (4, 8, 12):
    for chunk in (16, 64, 256):
        t = synth.init_tables(d.num_users, d.num_items, D)
        ctx = abi.Context(D, K, d.num_users, d.num_items, warps_per_block=warps, chunk=chunk)
        ctx.set_dataset(d.clicks, d.his_items, d.masks); ctx.bind_tables(*t)
        ms = []
        for ep in range(3):
            loss = ctx.train_epoch(); ms.append(ctx.last_kernel_ms())
        best = min(ms)
        print(f"warps {warps:2d} chunk {chunk:3d}  ms {ms}  loss {loss:.4f}  samples/s {d.train_size / best * 1e3:.3e}  GB/s {bytes_epoch / best / 1e6:.0f}")
        ctx.close()
