"""Generate tests/golden/*.npz from the UNMODIFIED reference.

Run in the build container (needs /root/reference and oracle/_ref built):
    python tools/make_golden.py
Training vectors come from oracle/_ref/libheat_ref.so (the reference's own C++ objects,
forward_backward / BehaviorAggregator / SGD, driven step by step); evaluation vectors from the
reference's Engine::evaluate0 and its Python cf/metrics.py imported from /root/reference.
Inputs are stored next to the outputs so the fixtures do not depend on numpy's RNG streams.
"""
from __future__ import annotations

import contextlib
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from heat_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
SEED = 2022


def train_case(name, U, I, N, D, K, H, clip, lr, steps_at=(1, 3, 40), negs_keep=None):
    d = synth.generate(U, I, N, max_his=H, seed=SEED, n_clusters=4, max_degree=I // 2)
    uw, iw, w0 = synth.init_tables(U, I, D, seed=SEED)
    init = dict(user_w=uw.copy(), item_w=iw.copy(), w0=w0.copy())
    negs = oracle.neg_table(SEED, 0, 0, d.clicks, K, I)
    if negs_keep is not None:          # large num_negs: keep the table small (tests use the first rows only)
        negs = negs[:negs_keep].copy()
    ref = oracle.RefOracle(D, K, d.clicks, d.his_items, d.masks, uw, iw, w0, clip_val=clip, l_r=lr)
    out = dict(clicks=d.clicks, his_items=d.his_items, masks=d.masks, negs=negs,
               D=D, K=K, clip=np.float32(clip), lr=np.float32(lr), seed=SEED, **{"init_" + k: v for k, v in init.items()})
    losses = []
    for s in range(max(steps_at)):
        u, p = map(int, d.clicks[s])
        losses.append(ref.step(u, p, negs[s]))
        if s + 1 in steps_at:
            gu, gi = ref.grads()
            t = f"s{s + 1}_"
            out.update({t + "user_w": uw.copy(), t + "item_w": iw.copy(), t + "w0": w0.copy(),
                        t + "user_g": gu, t + "item_g": gi})
    out["losses"] = np.array(losses, np.float32)
    ref.close()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, "losses[:3]", losses[:3])


def eval_case(name, U, I, N, D, H):
    sys.path.insert(0, "/root/reference/cf_cpu/cf")
    import metrics as ref_metrics  # the reference's own numpy metrics

    d = synth.generate(U, I, N, max_his=H, seed=SEED + 7, n_clusters=4, max_degree=I // 2)
    rng = np.random.default_rng(SEED + 9)
    uw = rng.standard_normal((U, D)).astype(np.float32)
    iw = rng.standard_normal((I, D)).astype(np.float32)
    w0 = np.zeros((D, D), np.float32)
    ref = oracle.RefOracle(D, 4, d.clicks, d.his_items, d.masks, uw, iw, w0)
    scores = ref.scores()
    ref.close()

    class _DS:  # the two attributes metrics.evaluate_metrics reads (metrics.py:8-10)
        def __init__(self, indptr, items):
            self.user_items_dic = {u: items[indptr[u]:indptr[u + 1]].tolist() for u in range(U)
                                   if indptr[u + 1] > indptr[u]}

    names = ['Recall(k=20)', 'Recall(k=50)', 'NDCG(k=20)', 'NDCG(k=50)', 'HitRate(k=20)', 'HitRate(k=50)']
    # metrics.py eval()s the metric strings in its own namespace
    with contextlib.redirect_stdout(io.StringIO()):
        res = ref_metrics.evaluate_metrics(_DS(d.train_indptr, d.train_items), _DS(d.test_indptr, d.test_items),
                                           scores.copy(), names)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), user_w=uw, item_w=iw, scores=scores,
                        train_indptr=d.train_indptr, train_items=d.train_items,
                        test_indptr=d.test_indptr, test_items=d.test_items,
                        metric_names=np.array(names), metric_values=np.array([res[n] for n in names]))
    print(name, res)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    assert oracle.RefOracle.available(), "build oracle/_ref first: make -C oracle ref"
    train_case("train_d64_k16", U=64, I=96, N=1500, D=64, K=16, H=100, clip=1.0, lr=1e-2)
    train_case("train_d128_k16", U=48, I=80, N=1200, D=128, K=16, H=100, clip=0.1, lr=1e-2)
    train_case("train_d64_k16_dups", U=64, I=24, N=600, D=64, K=16, H=20, clip=1.0, lr=1e-2)
    train_case("train_d32_k8", U=40, I=64, N=800, D=32, K=8, H=50, clip=1.0, lr=1e-2)
    # the large-negative regime (BASELINE.json configs[3] is emb_dim 128, num_negs 1000); I=1000 makes every
    # sample carry many duplicate negatives, the other two have few
    train_case("train_d128_k1000", U=16, I=1000, N=300, D=128, K=1000, H=100, clip=1.0, lr=1e-2, steps_at=(1, 3), negs_keep=72)
    train_case("train_d64_k200", U=32, I=2000, N=400, D=64, K=200, H=50, clip=1.0, lr=1e-2, steps_at=(1, 3), negs_keep=72)
    train_case("train_d32_k48", U=32, I=500, N=400, D=32, K=48, H=50, clip=0.1, lr=1e-2, steps_at=(1, 3), negs_keep=72)
    eval_case("eval_d64", U=96, I=400, N=3000, D=64, H=100)
