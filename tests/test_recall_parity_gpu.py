"""north_star's third correctness leg, driver-run: per-epoch loss and final Recall@20 against the reference C++ backend
on identical synthetic inputs (what cf_cpu/cf/main.py:115-124 + cf/metrics.py:5-36 report), for BASELINE.json configs[0]
(Gowalla shape, 1 epoch), [1] (Yelp18 shape, 5 epochs, k = 20/50) and [2] (Amazon-Books shape, 5 epochs) -- for the
library DEFAULTS (engine 0: the warp-per-sample kernel on its full grid) and for the loss-strict engine 2 (CTA per
sample, one sample in flight per ~600 item rows).

Reference arm = oracle/_ref (the unmodified reference sources compiled in place) when it travelled with the repo, else
the C restatement that is pinned to it (tests/test_oracle.py); all host threads; `sampler=2`, so both arms draw the same
Philox negatives.  The reference is a Hogwild program whose result moves with its thread schedule, so its own run-to-run
spread is MEASURED here (REF_RUNS runs per shape) and the stated tolerances are
    |mean GPU Recall@20 - mean reference Recall@20| <= 0.005 + reference spread + GPU spread (max - min over GPU_RUNS runs)
        (north_star: 0.005 absolute; both engines; both arms are Hogwild programs whose result moves by ~0.002 run to run)
    |loss_e - mean reference loss_e| <= LOSS_TOL + reference spread of that epoch, every epoch, where LOSS_TOL is
        0.03 for engine 2, and for the default engine 0.15 in the first epoch and 0.06 afterwards: with ~1 800 samples
        in flight the loss trails the reference's early on (+0.06...+0.13 in epoch 0, measured) and meets it by epoch 20
        (DESIGN.md section 6 has the 20-epoch trajectories).
"""
import os

import numpy as np
import pytest

import oracle
from heat_b200 import abi, synth

pytestmark = pytest.mark.gpu

D, K = 64, 16
RECALL_TOL = 0.005
GPU_RUNS = 3
LOSS_TOL = {2: (0.03, 0.03), 0: (0.15, 0.06)}         # engine -> (first epoch, later epochs)
THREADS = os.cpu_count() or 1
# shape, epochs (BASELINE.json configs[0..2]), clip_val of the shape's YAML, reference runs (spread)
CASES = [("gowalla", 1, 0.1, 3), ("yelp18", 5, 1.0, 3), ("amazon-book", 5, 1.0, 2)]


def _reference(d, epochs, clip):
    t = synth.init_tables(d.num_users, d.num_items, D)
    cls = oracle.RefOracle if oracle.RefOracle.available() else oracle.COracle
    o = cls(D, K, d.clicks, d.his_items, d.masks, *t, clip_val=clip)
    if cls is oracle.RefOracle:
        losses = [o.train_epoch(nthreads=THREADS, sampler=2) for _ in range(epochs)]
    else:
        losses = [o.train_epoch(nthreads=THREADS) for _ in range(epochs)]
    o.close()
    return losses, t, cls.__name__


def _evaluate(d, tables):
    ctx = abi.Context(D, K, d.num_users, d.num_items)
    ctx.bind_tables(*tables)
    res, _ = ctx.eval_metrics(d.train_indptr, d.train_items, d.test_indptr, d.test_items)
    ctx.close()
    return res


_REF_CACHE = {}


def _reference_runs(name, epochs, clip, ref_runs):
    """(dataset, per-run losses [runs, epochs], Recall@20 per run, Recall@50 per run, kind), computed once per shape"""
    if name not in _REF_CACHE:
        d = synth.generate_named(name)
        ref_losses, ref_recall, ref_recall50 = [], [], []
        kind = ""
        for _ in range(ref_runs):
            losses, tables, kind = _reference(d, epochs, clip)
            res = _evaluate(d, tables)
            ref_losses.append(losses); ref_recall.append(res["Recall(k=20)"]); ref_recall50.append(res["Recall(k=50)"])
        _REF_CACHE[name] = (d, np.array(ref_losses), ref_recall, ref_recall50, kind)
    return _REF_CACHE[name]


@pytest.mark.parametrize("engine", [0, 2])
@pytest.mark.parametrize("name,epochs,clip,ref_runs", CASES)
def test_epoch_losses_and_recall_match_the_reference_backend(name, epochs, clip, ref_runs, engine):
    d, ref_losses, ref_recall, ref_recall50, kind = _reference_runs(name, epochs, clip, ref_runs)
    loss_mean, loss_spread = ref_losses.mean(0), ref_losses.max(0) - ref_losses.min(0)
    r_mean, r_spread = float(np.mean(ref_recall)), float(np.max(ref_recall) - np.min(ref_recall))

    # the GPU arm is a Hogwild program too: GPU_RUNS runs, their mean against the reference's mean, their spread added
    runs = []
    for _ in range(GPU_RUNS):
        t = synth.init_tables(d.num_users, d.num_items, D)
        ctx = abi.Context(D, K, d.num_users, d.num_items, clip_val=clip, engine=engine)       # engine 0 = library defaults
        ctx.set_dataset(d.clicks, d.his_items, d.masks)
        ctx.bind_tables(*t)
        run_losses = [ctx.train_epoch() for _ in range(epochs)]
        ran = ctx.last_engine()
        assert ran == (1 if engine == 0 else 2)
        ctx.sync_to_host()
        ctx.close()
        runs.append((run_losses, _evaluate(d, t)))
    losses = np.mean([r[0] for r in runs], axis=0)
    g20 = [r[1]["Recall(k=20)"] for r in runs]
    g50 = [r[1]["Recall(k=50)"] for r in runs]
    res = {"Recall(k=20)": float(np.mean(g20)), "Recall(k=50)": float(np.mean(g50))}
    g_spread = float(np.max(g20) - np.min(g20))
    print(f"{name}: reference ({kind}, {THREADS} threads, {ref_runs} runs) loss {np.round(loss_mean, 4).tolist()} "
          f"spread {np.round(loss_spread, 4).tolist()} Recall@20 {r_mean:.4f} (spread {r_spread:.4f}) Recall@50 {np.mean(ref_recall50):.4f} | "
          f"GPU engine {engine} (kernel family {ran}, {GPU_RUNS} runs) loss {np.round(losses, 4).tolist()} Recall@20 {res['Recall(k=20)']:.4f} "
          f"({res['Recall(k=20)'] - r_mean:+.4f}, spread {g_spread:.4f}) Recall@50 {res['Recall(k=50)']:.4f}")
    for e in range(epochs):
        tol = LOSS_TOL[engine][0 if e == 0 else 1]
        assert abs(losses[e] - loss_mean[e]) <= tol + loss_spread[e], (e, losses[e], loss_mean[e], loss_spread[e])
    assert abs(res["Recall(k=20)"] - r_mean) <= RECALL_TOL + r_spread + g_spread, (res["Recall(k=20)"], r_mean, r_spread, g_spread)
    assert abs(res["Recall(k=50)"] - float(np.mean(ref_recall50))) <= 2 * RECALL_TOL + float(np.max(ref_recall50) - np.min(ref_recall50)) + float(np.max(g50) - np.min(g50))


def test_recall_of_reference_tables_agrees_with_the_reference_metric_code():
    """The Recall both arms are judged by comes from heat_eval_metrics.  On tables the REFERENCE trained, a sample of
    users is scored on the host (fp32 U.I^T as Engine::evaluate0, engine.cpp:388-400) and ranked by the restatement of
    cf/metrics.py; the fused GPU evaluation of the same users must give the same numbers."""
    d = synth.generate_named("gowalla")
    _, (uw, iw, w0), _ = _reference(d, 1, 0.1)
    users = np.random.default_rng(5).choice(d.num_users, 600, replace=False)
    users.sort()

    def sub(indptr, items):
        rows = [items[indptr[u]:indptr[u + 1]] for u in users]
        return np.concatenate([[0], np.cumsum([r.size for r in rows])]).astype(np.int64), np.concatenate(rows).astype(np.int32)
    tr_p, tr_i = sub(d.train_indptr, d.train_items)
    te_p, te_i = sub(d.test_indptr, d.test_items)
    su = np.ascontiguousarray(uw[users])
    _, _, want = oracle.topk_and_metrics(su @ iw.T, tr_p, tr_i, te_p, te_i)
    ctx = abi.Context(D, K, users.size, d.num_items)
    ctx.bind_tables(su, iw, w0)
    got, _ = ctx.eval_metrics(tr_p, tr_i, te_p, te_i)
    ctx.close()
    for k, v in want.items():
        assert abs(got[k] - v) < 2e-3, (k, got[k], v)      # a near-tie at rank k may swap between fp32 summation orders
