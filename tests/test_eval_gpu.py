"""GPU parity tests of the evaluation path (tcgen05 score GEMM + fused masked top-k + metrics),
through the C ABI, against the golden vectors of the compiled reference / its metrics.py and the
oracle."""
import numpy as np
import pytest

import oracle
from heat_b200 import abi, synth
from helpers import load_golden, rel_err

pytestmark = pytest.mark.gpu


def ctx_for(user_w, item_w):
    U, D = user_w.shape
    ctx = abi.Context(D, 4, U, item_w.shape[0])
    ctx.bind_tables(user_w, item_w, np.zeros((D, D), np.float32))
    return ctx


def check_topk(ids, sc, scores, train_indptr, train_items, k, tie_rel=1e-6):
    """ids/sc from the device vs an fp64 restatement on the given score matrix: same ids except
    where neighbouring scores are closer than fp32 summation noise."""
    U, I = scores.shape
    for u in range(U):
        row = scores[u].astype(np.float64).copy()
        if train_indptr is not None:
            row[train_items[train_indptr[u]:train_indptr[u + 1]]] = -np.inf
        order = np.lexsort((np.arange(I), -row))[:k]
        n_valid = int(np.isfinite(row[order]).sum())
        got = ids[u]
        assert (got[:n_valid] >= 0).all() and (got[n_valid:] == -1).all(), (u, got, n_valid)
        assert len(set(got[:n_valid].tolist())) == n_valid
        # scores returned are the scores of the ids returned, descending
        assert np.allclose(sc[u, :n_valid], scores[u, got[:n_valid]], rtol=1e-5, atol=1e-6), u
        assert (np.diff(sc[u, :n_valid]) <= 0).all(), u
        scale = max(1.0, float(np.abs(row[order[:n_valid]]).max())) if n_valid else 1.0
        for r in range(n_valid):
            if got[r] != order[r]:
                assert abs(row[got[r]] - row[order[r]]) <= tie_rel * scale * 10, (u, r, got[r], order[r])
        if train_indptr is not None:
            assert not set(got[:n_valid].tolist()) & set(train_items[train_indptr[u]:train_indptr[u + 1]].tolist())


def test_dense_scores_match_reference_golden():
    g = load_golden("eval_d64")
    ctx = ctx_for(g["user_w"].copy(), g["item_w"].copy())
    sc = ctx.eval_dense()
    assert sc.shape == g["scores"].shape
    assert rel_err(sc, g["scores"]) < 1e-5, rel_err(sc, g["scores"])


def test_topk_and_metrics_match_reference_golden():
    g = load_golden("eval_d64")
    ctx = ctx_for(g["user_w"].copy(), g["item_w"].copy())
    ids, sc = ctx.eval_topk(50, g["train_indptr"], g["train_items"])
    check_topk(ids, sc, g["scores"], g["train_indptr"], g["train_items"], 50)
    oids, _, _ = oracle.topk_and_metrics(g["scores"], g["train_indptr"], g["train_items"], g["test_indptr"], g["test_items"])
    assert (ids == oids).mean() > 0.999
    res, sums = ctx.eval_metrics(g["train_indptr"], g["train_items"], g["test_indptr"], g["test_items"])
    for n, v in zip(g["metric_names"], g["metric_values"]):      # values from the reference's cf/metrics.py
        assert abs(res[str(n)] - float(v)) < 1e-6, (n, res[str(n)], float(v))
    assert sums[6] == np.count_nonzero(np.diff(g["test_indptr"]))
    assert ctx.last_eval_ms() > 0


@pytest.mark.parametrize("D,U,I", [(64, 300, 1000), (128, 131, 517), (32, 257, 90), (64, 128, 64), (64, 1, 1)])
def test_dense_and_topk_shapes(D, U, I):
    rng = np.random.default_rng(D + U + I)
    uw = rng.standard_normal((U, D)).astype(np.float32)
    iw = rng.standard_normal((I, D)).astype(np.float32)
    iw[I // 2] = iw[0]                                   # an exact tie: the smaller id must win
    ref = uw.astype(np.float64) @ iw.astype(np.float64).T
    ctx = ctx_for(uw, iw)
    sc = ctx.eval_dense()
    assert rel_err(sc, ref) < 2e-6, rel_err(sc, ref)     # 3xTF32: fp32-level, not TF32-level (1e-3)
    k = 20
    ids, s = ctx.eval_topk(k)
    check_topk(ids, s, sc, None, None, min(k, I) if I >= k else k)
    if I // 2 > 0:
        for u in range(U):
            pos0, pos1 = np.where(ids[u] == 0)[0], np.where(ids[u] == I // 2)[0]
            if len(pos0) and len(pos1):
                assert pos0[0] < pos1[0]


def test_topk_with_train_mask_medium():
    d = synth.generate(1500, 4000, 60000, max_his=50, seed=5)
    uw, iw, _ = synth.init_tables(d.num_users, d.num_items, 64, seed=6)
    uw *= 30; iw *= 30
    ctx = ctx_for(uw, iw)
    sc = ctx.eval_dense()
    ids, s = ctx.eval_topk(50, d.train_indptr, d.train_items)
    check_topk(ids, s, sc, d.train_indptr, d.train_items, 50)
    res, sums = ctx.eval_metrics(d.train_indptr, d.train_items, d.test_indptr, d.test_items)
    _, _, want = oracle.topk_and_metrics(sc, d.train_indptr, d.train_items, d.test_indptr, d.test_items)
    for n in want:
        assert abs(res[n] - want[n]) < 1e-6, (n, res[n], want[n])


def test_eval_follows_training():
    """evaluate after training uses the CURRENT device tables (raw stored user rows, engine.cpp:392-398)."""
    d = synth.generate_named("small")
    t = synth.init_tables(d.num_users, d.num_items, 64)
    ctx = abi.Context(64, 16, d.num_users, d.num_items, max_blocks=8, warps_per_block=4)
    ctx.set_dataset(d.clicks, d.his_items, d.masks); ctx.bind_tables(*t)
    r0, _ = ctx.eval_metrics(d.train_indptr, d.train_items, d.test_indptr, d.test_items)
    for _ in range(4):
        ctx.train_epoch()
    r1, _ = ctx.eval_metrics(d.train_indptr, d.train_items, d.test_indptr, d.test_items)
    ctx.sync_to_host()
    _, _, want = oracle.topk_and_metrics(t[0] @ t[1].T, d.train_indptr, d.train_items, d.test_indptr, d.test_items)
    print("recall@20 before/after", r0["Recall(k=20)"], r1["Recall(k=20)"], "host restatement", want["Recall(k=20)"])
    assert r1["Recall(k=20)"] > 2 * r0["Recall(k=20)"]
    assert abs(r1["Recall(k=20)"] - want["Recall(k=20)"]) < 2e-3


def test_full_size_amazon_shape_eval_properties():
    d = synth.generate_named("amazon-book")
    uw, iw, _ = synth.init_tables(d.num_users, d.num_items, 64, seed=9)
    ctx = ctx_for(uw, iw)
    ids, sc = ctx.eval_topk(50, d.train_indptr, d.train_items)
    print("amazon-shape fused GEMM+top-50 ms", ctx.last_eval_ms())
    assert (ids >= 0).all() and (ids < d.num_items).all()
    assert (np.diff(sc, axis=1) <= 0).all()
    rng = np.random.default_rng(1)
    for u in rng.integers(0, d.num_users, 40):
        row = (uw[u].astype(np.float64) @ iw.astype(np.float64).T)
        row[d.train_items[d.train_indptr[u]:d.train_indptr[u + 1]]] = -np.inf
        want = np.lexsort((np.arange(d.num_items), -row))[:50]
        same = (ids[u] == want)
        assert same.mean() > 0.9 and np.allclose(sc[u], row[ids[u]], rtol=1e-5, atol=1e-7), (u, same.mean())
        assert set(ids[u].tolist()) == set(want.tolist()) or np.abs(row[ids[u]].min() - row[want].min()) < 1e-9


def test_popularity_ordered_streaming_does_not_change_the_result():
    """The fused top-k streams the items in order of descending train-set degree (eval_kernel.cu); ids, scores, masking and ties must
    be those of the id-ordered definition: a table whose row norms spread over 6x, exact duplicates of rows (ties) and a
    train mask."""
    rng = np.random.default_rng(11)
    U, I, D = 300, 5000, 64
    uw = (rng.standard_normal((U, D)) * 0.1).astype(np.float32)
    iw = (rng.standard_normal((I, D)) * 0.1).astype(np.float32) * rng.uniform(0.5, 3.0, (I, 1)).astype(np.float32)
    iw[100:140] = iw[4000:4040]                                        # exact ties between far-apart ids
    deg = rng.integers(0, 40, U)
    indptr = np.concatenate([[0], np.cumsum(deg)]).astype(np.int64)
    items = np.concatenate([np.sort(rng.choice(I, d_, replace=False)) for d_ in deg]).astype(np.int32)
    ctx = ctx_for(uw, iw)
    ids, sc = ctx.eval_topk(50, indptr, items)
    ctx.close()
    check_topk(ids, sc, uw.astype(np.float64) @ iw.astype(np.float64).T, indptr, items, 50)


def test_single_pass_tf32_precision_knob():
    """heat_config.eval_precision = 1: one TF32 pass instead of three.  Reported: top-50 set overlap and the change of
    Recall@20 against the default (fp32-level) path on the reference-generated evaluation fixture."""
    g = load_golden("eval_d64")
    U, D = g["user_w"].shape
    out = {}
    for prec in (0, 1):
        ctx = abi.Context(D, 4, U, g["item_w"].shape[0], eval_precision=prec)
        ctx.bind_tables(g["user_w"].copy(), g["item_w"].copy(), np.zeros((D, D), np.float32))
        ids, _ = ctx.eval_topk(50, g["train_indptr"], g["train_items"])
        res, _ = ctx.eval_metrics(g["train_indptr"], g["train_items"], g["test_indptr"], g["test_items"])
        ctx.close()
        out[prec] = (ids, res)
    overlap = np.mean([len(set(a.tolist()) & set(b.tolist())) / 50.0 for a, b in zip(out[0][0], out[1][0])])
    d20 = out[1][1]["Recall(k=20)"] - out[0][1]["Recall(k=20)"]
    print(f"single-pass TF32: top-50 set overlap {overlap:.4f}, dRecall@20 {d20:+.5f}")
    assert overlap > 0.97 and abs(d20) < 3e-3
