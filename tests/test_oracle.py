"""CPU tests: the oracle against known-answer vectors, the golden fixtures generated from the
compiled reference, and (when oracle/_ref exists) the compiled reference itself."""
import numpy as np
import pytest

import oracle
from heat_b200 import synth
from helpers import BIGK_CASES, TRAIN_CASES, fresh_tables, load_golden, rel_err


def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32-10
    assert oracle.philox4x32_10([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert oracle.philox4x32_10([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert oracle.philox4x32_10([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344],
                                [0xa4093822, 0x299f31d0]) == [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


@pytest.mark.parametrize("name", TRAIN_CASES + BIGK_CASES)
def test_neg_table_matches_golden_and_rules(name):
    g = load_golden(name)
    K, I = int(g["K"]), g["init_item_w"].shape[0]
    negs = oracle.neg_table(int(g["seed"]), 0, 0, g["clicks"], K, I)
    assert np.array_equal(negs[:g["negs"].shape[0]], g["negs"])
    assert negs.max() < I
    # look-back rule == the reference's persistent vector (uniform_random_negative_sampler.cpp:26-36)
    persistent = np.zeros(K, np.uint64)
    for s in range(min(300 if K <= 16 else 12, g["clicks"].shape[0])):
        for k in range(K):
            c = [s & 0xffffffff, s >> 32, 0, k >> 2]
            r = oracle.philox4x32_10(c, [int(g["seed"]) & 0xffffffff, int(g["seed"]) >> 32])[k & 3]
            d = (r * I) >> 32
            if d != int(g["clicks"][s, 1]):
                persistent[k] = d
        assert np.array_equal(persistent, negs[s]), s


@pytest.mark.parametrize("name", TRAIN_CASES + BIGK_CASES)
@pytest.mark.parametrize("steps", [1, 3])
def test_c_oracle_matches_reference_golden(name, steps):
    g = load_golden(name)
    uw, iw, w0 = fresh_tables(g)
    o = oracle.COracle(int(g["D"]), int(g["K"]), g["clicks"], g["his_items"], g["masks"], uw, iw, w0,
                       clip_val=float(g["clip"]), l_r=float(g["lr"]))
    losses = [o.step(int(g["clicks"][s, 0]), int(g["clicks"][s, 1]), g["negs"][s]) for s in range(steps)]
    gu, gi = o.grads()
    t = f"s{steps}_"
    tol = 1e-5 if steps == 1 else 2e-4     # the dynamics amplify rounding differences step by step
    assert np.allclose(losses, g["losses"][:steps], rtol=tol * 10, atol=1e-5)
    for got, key in ((uw, "user_w"), (iw, "item_w"), (w0, "w0"), (gu, "user_g"), (gi, "item_g")):
        # weights: the north-star's 1e-5; raw grad rows carry the cancellation in (uu*p - up*u) and get 10x
        lim = tol * (10 if key.endswith("_g") else 1)
        assert rel_err(got, g[t + key]) < lim, (key, rel_err(got, g[t + key]))


def test_c_oracle_w0_update_matches_golden_when_tables_frozen():
    # 40 steps include one W0 update (iteration 32); with the table LR at 0 nothing is chaotic.
    g = load_golden("train_d64_k16")
    if not oracle.RefOracle.available():
        pytest.skip("oracle/_ref not built")
    a, b = fresh_tables(g), fresh_tables(g)
    co = oracle.COracle(64, 16, g["clicks"], g["his_items"], g["masks"], *a)
    ro = oracle.RefOracle(64, 16, g["clicks"], g["his_items"], g["masks"], *b)
    co.set_lr(0.0); ro.set_lr(0.0)
    for s in range(70):
        u, p = map(int, g["clicks"][s])
        l1, l2 = co.step(u, p, g["negs"][s]), ro.step(u, p, g["negs"][s])
        assert abs(l1 - l2) < 1e-4 * max(1.0, abs(l2))
    assert not np.array_equal(a[2], g["init_w0"])          # W0 did move
    assert rel_err(a[2], b[2]) < 1e-5
    assert not np.array_equal(a[0], g["init_user_w"])     # user rows are overwritten with the aggregate even at lr 0
    assert rel_err(a[0], b[0]) < 1e-5 and rel_err(a[1], b[1]) < 1e-5


@pytest.mark.skipif(not oracle.RefOracle.available(), reason="oracle/_ref not built (no /root/reference)")
def test_c_oracle_epoch_statistics_match_compiled_reference():
    d = synth.generate_named("tiny")
    D, K = 64, 16
    a = synth.init_tables(d.num_users, d.num_items, D); b = synth.init_tables(d.num_users, d.num_items, D)
    co = oracle.COracle(D, K, d.clicks, d.his_items, d.masks, *a)
    ro = oracle.RefOracle(D, K, d.clicks, d.his_items, d.masks, *b)
    for ep in range(3):
        l1 = co.train_epoch(nthreads=1)
        l2 = ro.train_epoch(nthreads=1, sampler=1)   # reference's own ignore_pos logic on the Philox draws
        assert abs(l1 - l2) / l2 < 0.05, (ep, l1, l2)
    gu, gi = co.grads()
    assert not gu.any() and not gi.any()              # zero_grad at epoch end (engine.cpp:345-347)


@pytest.mark.skipif(not oracle.RefOracle.available(), reason="oracle/_ref not built (no /root/reference)")
def test_reference_stateless_and_persistent_samplers_agree():
    d = synth.generate_named("tiny")
    D, K = 64, 16
    a = synth.init_tables(d.num_users, d.num_items, D); b = synth.init_tables(d.num_users, d.num_items, D)
    r1 = oracle.RefOracle(D, K, d.clicks, d.his_items, d.masks, *a)
    r2 = oracle.RefOracle(D, K, d.clicks, d.his_items, d.masks, *b)
    assert r1.train_epoch(nthreads=1, sampler=1) == r2.train_epoch(nthreads=1, sampler=2)
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[0], b[0])


def test_lr_schedule_single_milestone():
    d = synth.generate(40, 60, 400, max_his=10, n_clusters=2, max_degree=20)
    t = synth.init_tables(40, 60, 32)
    o = oracle.COracle(32, 8, d.clicks, d.his_items, d.masks, *t, milestones=(2,))
    lrs = []
    for _ in range(5):
        o.train_epoch(); lrs.append(o.get_lr())
    # optimizer.cpp:24-30: decay at epochs 2 and 4 (epoch > 0 and epoch % 2 == 0)
    assert np.allclose(lrs, [1e-2, 1e-2, 1e-3, 1e-3, 1e-4], rtol=1e-6)


def test_eval_oracle_matches_reference_metrics_golden():
    g = load_golden("eval_d64")
    U, D = g["user_w"].shape
    I = g["item_w"].shape[0]
    clicks = np.zeros((1, 2), np.uint64); his = np.zeros((U, 1), np.uint64); masks = np.ones((U, 1), np.uint64)
    o = oracle.COracle(D, 4, clicks, his, masks, g["user_w"].copy(), g["item_w"].copy(), np.zeros((D, D), np.float32))
    sc = o.scores()
    assert rel_err(sc, g["scores"]) < 1e-5
    _, _, res = oracle.topk_and_metrics(g["scores"], g["train_indptr"], g["train_items"],
                                        g["test_indptr"], g["test_items"])
    for n, v in zip(g["metric_names"], g["metric_values"]):
        assert abs(res[str(n)] - float(v)) < 1e-9, (n, res[str(n)], v)
