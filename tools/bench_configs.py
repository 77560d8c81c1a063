"""Per-config rows of BASELINE.json: the configs that are not the bench.py headline.

    python tools/bench_configs.py [gowalla yelp18 amazon-book-k1000 ...] > gpurun_out/configs.json

One JSON line per config: epoch ms (CUDA events around heat_train_epoch, L2 flushed between epochs), samples/s,
algorithmic GB/s of the training kernel (bench.algorithmic_bytes), per-epoch losses and -- with --eval -- the
evaluation time and Recall@20.  Resident inputs; same synthetic generator and seed as bench.py.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (shape, D, K, clip, epochs, BASELINE.json configs[] index)
    "gowalla": ("gowalla", 64, 16, 0.1, 1, 0),
    "yelp18": ("yelp18", 64, 16, 1.0, 5, 1),
    "amazon-book": ("amazon-book", 64, 16, 1.0, 3, 2),
    "amazon-book-k1000": ("amazon-book", 128, 1000, 1.0, 2, 3),
}


def main():
    import torch
    import bench
    from heat_b200 import abi, synth

    ap = argparse.ArgumentParser()
    ap.add_argument("configs", nargs="*", default=["gowalla", "yelp18", "amazon-book-k1000"])
    ap.add_argument("--eval", action="store_true")
    ap.add_argument("--engine", type=int, default=0)
    ap.add_argument("--max-blocks", type=int, default=0)
    ap.add_argument("--chunk", type=int, default=0)
    ap.add_argument("--samples", type=int, default=0, help="train on the first N samples only (0 = all)")
    ap.add_argument("--epochs", type=int, default=0)
    args = ap.parse_args()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for name in args.configs:
        shape, D, K, clip, epochs, idx = CONFIGS[name]
        epochs = args.epochs or epochs
        d = synth.generate_named(shape, max_his=100, seed=2022)
        clicks = d.clicks[:args.samples] if args.samples else d.clicks
        uw, iw, w0 = synth.init_tables(d.num_users, d.num_items, D, seed=2022)
        ctx = abi.Context(D, K, d.num_users, d.num_items, clip_val=clip, l_r=1e-2, milestones=(10,), seed=2022,
                          engine=args.engine, max_blocks=args.max_blocks, chunk=args.chunk)
        ctx.set_dataset(clicks, d.his_items, d.masks)
        ctx.bind_tables(uw, iw, w0)
        alg = bench.algorithmic_bytes(clicks, d.masks, D, K)
        ms, kms, losses = [], [], []
        for _ in range(epochs):
            flush.add_(1); torch.cuda.synchronize()
            losses.append(ctx.train_epoch())
            ms.append(ctx.last_epoch_ms()); kms.append(ctx.last_kernel_ms())
        best = float(np.min(kms))
        row = {"config": name, "baseline_configs_index": idx, "shape": shape, "users": d.num_users, "items": d.num_items,
               "samples": int(clicks.shape[0]), "emb_dim": D, "num_negs": K, "clip": clip, "epoch_ms": ms, "kernel_ms": kms,
               "samples_per_s": clicks.shape[0] / (float(np.mean(ms)) * 1e-3), "algorithmic_bytes_per_epoch": alg,
               "algorithmic_GBps": alg / (float(np.mean(kms)) * 1e-3) / 1e9, "best_algorithmic_GBps": alg / (best * 1e-3) / 1e9,
               "losses": losses}
        if args.eval:
            t0 = time.perf_counter()
            res, _ = ctx.eval_metrics(d.train_indptr, d.train_items, d.test_indptr, d.test_items)
            row["eval_wall_ms"] = (time.perf_counter() - t0) * 1e3
            row["eval_kernel_ms"] = ctx.last_eval_ms()
            row["metrics"] = res
        ctx.close()
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
