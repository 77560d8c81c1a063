"""Run-to-run spread of the compiled reference itself (oracle/_ref): final Recall@20 after `epochs` epochs at
several OpenMP thread counts, same seeded inputs and same Philox negatives.  CPU only.
usage: ref_spread.py <shape> <epochs> <threads...>"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from heat_b200 import synth

name, epochs = sys.argv[1], int(sys.argv[2])
threads = [int(x) for x in sys.argv[3:]] or [1, 4, 16]
clip = 0.1 if name == "gowalla" else 1.0
d = synth.generate_named(name)
out = []


def recall20(uw, iw, slab=2048):
    """cf/metrics.py Recall(k=20): hits / #test items, mean over users with test items; train items masked."""
    tot, n = 0.0, 0
    for u0 in range(0, d.num_users, slab):
        u1 = min(d.num_users, u0 + slab)
        sc = uw[u0:u1] @ iw.T
        for u in range(u0, u1):
            sc[u - u0, d.train_items[d.train_indptr[u]:d.train_indptr[u + 1]]] = -np.inf
        top = np.argpartition(-sc, 20, axis=1)[:, :20]
        for u in range(u0, u1):
            te = d.test_items[d.test_indptr[u]:d.test_indptr[u + 1]]
            if len(te):
                tot += len(np.intersect1d(top[u - u0], te)) / len(te)
                n += 1
    return tot / n


for nt in threads:
    t = synth.init_tables(d.num_users, d.num_items, 64)
    o = oracle.RefOracle(64, 16, d.clicks, d.his_items, d.masks, *t, clip_val=clip)
    losses = [o.train_epoch(nthreads=nt, sampler=2) for _ in range(epochs)]
    res = {"Recall(k=20)": recall20(t[0], t[1])}
    out.append({"threads": nt, "losses": losses, **res})
    print(nt, [round(x, 4) for x in losses], {k: round(float(v), 4) for k, v in res.items()}, flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump({"shape": name, "epochs": epochs, "runs": out}, open(f"gpurun_out/ref_spread_{name}.json", "w"), indent=1)
