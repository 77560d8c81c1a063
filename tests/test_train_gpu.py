"""GPU parity tests of the training path, through the C ABI, against the oracle / golden vectors."""
import numpy as np
import pytest

import oracle
from heat_b200 import abi, synth
from helpers import BIGK_CASES, TRAIN_CASES, fresh_tables, load_golden, rel_err

pytestmark = pytest.mark.gpu


def make_ctx(g, **kw):
    U, D = g["init_user_w"].shape
    I = g["init_item_w"].shape[0]
    ctx = abi.Context(D, int(g["K"]), U, I, clip_val=float(g["clip"]), l_r=float(g["lr"]), seed=int(g["seed"]), **kw)
    ctx.set_dataset(g["clicks"], g["his_items"], g["masks"])
    return ctx


@pytest.mark.parametrize("name", TRAIN_CASES + BIGK_CASES)
@pytest.mark.parametrize("epoch,offset", [(0, 0), (3, 0), (1, 123456789012)])
def test_sampler_bit_exact(name, epoch, offset):
    g = load_golden(name)
    I = g["init_item_w"].shape[0]
    ctx = make_ctx(g, sample_offset=offset)
    N = g["clicks"].shape[0]
    dev = ctx.sample_negatives(epoch, 0, N)
    ref = oracle.neg_table(int(g["seed"]), epoch, offset, g["clicks"], int(g["K"]), I)
    assert np.array_equal(dev, ref)
    if epoch == 0 and offset == 0:
        assert np.array_equal(dev[:g["negs"].shape[0]], g["negs"])
    part = ctx.sample_negatives(epoch, 17, 101)
    assert np.array_equal(part, ref[17:101])


@pytest.mark.parametrize("name", TRAIN_CASES)
def test_history_gather_indices_bit_exact(name):
    g = load_golden(name)
    ctx = make_ctx(g)
    N = g["clicks"].shape[0]
    ids, cnt = ctx.history_indices(0, N)
    users = g["clicks"][:, 0].astype(np.int64)
    H = g["masks"][users, 0]
    assert np.array_equal(cnt, H)
    want = g["his_items"][users].copy()
    want[np.arange(want.shape[1])[None, :] >= H[:, None]] = np.uint64(0xFFFFFFFFFFFFFFFF)
    assert np.array_equal(ids, want)


# engine 0 = the default for the shape (large num_negs -> train_kernel_bigk.cuh); engine 3 forces the
# large-negative kernel onto the small shapes too
ENGINE_CASES = [(n, 0) for n in TRAIN_CASES + BIGK_CASES] + [(n, 3) for n in TRAIN_CASES] + [(n, 1) for n in TRAIN_CASES]


@pytest.mark.parametrize("name,engine", ENGINE_CASES)
@pytest.mark.parametrize("steps", [1, 3])
def test_steps_match_reference_golden(name, engine, steps):
    g = load_golden(name)
    uw, iw, w0 = fresh_tables(g)
    ctx = make_ctx(g, deterministic=True, w0_tail_mode=1, engine=engine)
    ctx.bind_tables(uw, iw, w0)
    total, per = ctx.train_range(0, steps, per_sample=True)
    ctx.sync_to_host()
    gu, gi = ctx.grads()
    t = f"s{steps}_"
    tol = 1e-5 if steps == 1 else 2e-4
    assert np.allclose(per, g["losses"][:steps], rtol=10 * tol, atol=1e-5), (per, g["losses"][:steps])
    assert abs(total - float(np.sum(g["losses"][:steps].astype(np.float64)))) < 1e-4 * steps
    for got, key in ((uw, "user_w"), (iw, "item_w"), (w0, "w0"), (gu, "user_g"), (gi, "item_g")):
        lim = tol * (10 if key.endswith("_g") else 1)
        assert rel_err(got, g[t + key]) < lim, (key, rel_err(got, g[t + key]))


@pytest.mark.parametrize("name,engine", ENGINE_CASES)
def test_each_sample_alone_matches_oracle(name, engine):
    """Teacher-forced: every one of the first 64 samples, applied alone to the SAME initial state,
    must match the oracle's single step (covers duplicate negatives, draws equal to the positive,
    short and full histories)."""
    g = load_golden(name)
    D, K = int(g["D"]), int(g["K"])
    ctx = make_ctx(g, deterministic=True, w0_tail_mode=1, engine=engine)
    uw, iw, w0 = fresh_tables(g)
    ctx.bind_tables(uw, iw, w0)
    worst = 0.0
    for s in range(64):
        uw[:], iw[:], w0[:] = fresh_tables(g)
        ctx.upload_tables(); ctx.zero_grad()
        _, per = ctx.train_range(s, s + 1, per_sample=True)
        ctx.sync_to_host()
        a = fresh_tables(g)
        o = oracle.COracle(D, K, g["clicks"], g["his_items"], g["masks"], *a, clip_val=float(g["clip"]), l_r=float(g["lr"]))
        l = o.step(int(g["clicks"][s, 0]), int(g["clicks"][s, 1]), g["negs"][s])
        o.close()
        assert abs(per[0] - l) < 1e-4 * max(1.0, abs(l)), (s, per[0], l)
        for got, ref in ((uw, a[0]), (iw, a[1])):
            worst = max(worst, rel_err(got, ref))
            assert rel_err(got, ref) < 1e-5, (s, rel_err(got, ref))
    print(name, "worst single-step rel err", worst)


@pytest.mark.parametrize("name,engine", [("train_d64_k16", 0), ("train_d128_k16", 0), ("train_d32_k8", 0), ("train_d64_k16", 3),
                                         ("train_d128_k16", 3), ("train_d128_k1000", 0), ("train_d64_k200", 0), ("train_d32_k48", 0)])
def test_w0_update_with_frozen_tables(name, engine):
    """70 sequential samples with the table LR at 0: two W0 updates (iterations 32 and 64), the tail
    dropped as the reference does; nothing chaotic, so the whole state must agree."""
    g = load_golden(name)
    D, K = int(g["D"]), int(g["K"])
    uw, iw, w0 = fresh_tables(g)
    ctx = make_ctx(g, deterministic=True, w0_tail_mode=1, engine=engine)
    ctx.bind_tables(uw, iw, w0)
    ctx.set_lr(0.0)
    _, per = ctx.train_range(0, 70, per_sample=True)
    ctx.sync_to_host()
    a = fresh_tables(g)
    o = oracle.COracle(D, K, g["clicks"], g["his_items"], g["masks"], *a, clip_val=float(g["clip"]), l_r=float(g["lr"]))
    o.set_lr(0.0)
    ref = [o.step(int(g["clicks"][s, 0]), int(g["clicks"][s, 1]), g["negs"][s]) for s in range(70)]
    assert np.allclose(per, ref, rtol=1e-4, atol=1e-5)
    assert not np.array_equal(w0, g["init_w0"])
    assert rel_err(w0, a[2]) < 1e-5, rel_err(w0, a[2])
    assert rel_err(uw, a[0]) < 1e-5 and rel_err(iw, a[1]) < 1e-5


def test_w0_tail_flush_applies_leftover():
    g = load_golden("train_d64_k16")
    outs = []
    for mode in (0, 1):
        uw, iw, w0 = fresh_tables(g)
        ctx = make_ctx(g, deterministic=True, w0_tail_mode=mode)
        ctx.bind_tables(uw, iw, w0); ctx.set_lr(0.0)
        ctx.train_range(0, 20); ctx.sync_to_host()
        outs.append(w0.copy())
    assert np.array_equal(outs[1], g["init_w0"])        # dropped: fewer than 32 samples
    assert not np.array_equal(outs[0], g["init_w0"])    # default: applied with the 1/32 scale


def test_epoch_statistics_hogwild_vs_sequential_oracle():
    d = synth.generate_named("small")
    D, K = 64, 16
    a = synth.init_tables(d.num_users, d.num_items, D)
    b = synth.init_tables(d.num_users, d.num_items, D)
    # 8 warps in flight ~ the reference at 8 OpenMP threads: on a 3000-item table a full grid
    # (1776 warps) would have every row in flight several times over, which no CPU run resembles
    ctx = abi.Context(D, K, d.num_users, d.num_items, max_blocks=2, warps_per_block=4)
    ctx.set_dataset(d.clicks, d.his_items, d.masks); ctx.bind_tables(*a)
    o = oracle.COracle(D, K, d.clicks, d.his_items, d.masks, *b)
    for ep in range(3):
        lg, lo = ctx.train_epoch(), o.train_epoch(nthreads=1)
        print("epoch", ep, "gpu", lg, "oracle", lo)
        assert np.isfinite(lg) and abs(lg - lo) / lo < 0.08, (ep, lg, lo)
    ctx.sync_to_host()
    gu, gi = ctx.grads()
    assert not gu.any() and not gi.any()
    assert ctx.get_epoch() == 3
    for t in a:
        assert np.isfinite(t).all()


def test_deterministic_mode_is_reproducible_and_hogwild_is_close():
    d = synth.generate_named("tiny")
    D, K = 64, 16
    runs = []
    for _ in range(2):
        t = synth.init_tables(d.num_users, d.num_items, D)
        ctx = abi.Context(D, K, d.num_users, d.num_items, deterministic=True)
        ctx.set_dataset(d.clicks, d.his_items, d.masks); ctx.bind_tables(*t)
        runs.append((ctx.train_epoch(), ctx.train_epoch()))
        ctx.sync_to_host(); runs.append(t[1].copy())
    assert runs[0] == runs[2] and np.array_equal(runs[1], runs[3])


def test_lr_schedule_and_bookkeeping():
    d = synth.generate(40, 60, 400, max_his=10, n_clusters=2, max_degree=20)
    t = synth.init_tables(40, 60, 32)
    ctx = abi.Context(32, 8, 40, 60, milestones=(2,))
    ctx.set_dataset(d.clicks, d.his_items, d.masks); ctx.bind_tables(*t)
    lrs = []
    for _ in range(5):
        ctx.train_epoch(); lrs.append(ctx.get_lr())
    assert np.allclose(lrs, [1e-2, 1e-2, 1e-3, 1e-3, 1e-4], rtol=1e-6)
    assert ctx.launch_count() >= 5


def test_errors_are_loud():
    ctx = abi.Context(64, 16, 10, 10)
    with pytest.raises(abi.HeatError):
        ctx.train_epoch()                                  # no dataset
    bad = np.array([[0, 11]], np.uint64)                   # item id out of range
    with pytest.raises(abi.HeatError, match="out of range"):
        ctx.set_dataset(bad, np.zeros((10, 4), np.uint64), np.ones((10, 1), np.uint64))
    with pytest.raises(abi.HeatError, match="unsupported"):
        c2 = abi.Context(48, 16, 10, 10)
        c2.set_dataset(np.array([[0, 1]], np.uint64), np.zeros((10, 4), np.uint64), np.ones((10, 1), np.uint64))
        c2.bind_tables(np.zeros((10, 48), np.float32), np.zeros((10, 48), np.float32), np.zeros((48, 48), np.float32))
        c2.train_epoch()


def test_full_size_amazon_shape_properties():
    """BASELINE configs[2] shape: one Hogwild epoch, size-independent properties only."""
    d = synth.generate_named("amazon-book")
    D, K = 64, 16
    t = synth.init_tables(d.num_users, d.num_items, D)
    ctx = abi.Context(D, K, d.num_users, d.num_items)
    ctx.set_dataset(d.clicks, d.his_items, d.masks); ctx.bind_tables(*t)
    # index contracts at full size on a strided sample
    negs = ctx.sample_negatives(0, 1000000, 1000512)
    ref = oracle.neg_table(2022, 0, 0, d.clicks, K, d.num_items)[1000000:1000512] if False else None
    sub = np.ascontiguousarray(d.clicks[:1000512])
    assert np.array_equal(negs, oracle.neg_table(2022, 0, 0, sub, K, d.num_items)[1000000:])
    l0 = ctx.train_epoch(); ms = ctx.last_kernel_ms()
    l1 = ctx.train_epoch()
    print("amazon-shape epoch loss", l0, l1, "kernel ms", ms, "samples/s", d.train_size / ms * 1e3)
    assert np.isfinite(l0) and np.isfinite(l1) and l1 < l0 < np.log(K + 1) * 1.5
    ctx.sync_to_host()
    assert all(np.isfinite(x).all() for x in t)
    assert ms < 2000.0     # the north-star's < 2 s epoch


def test_checkpoint_resume_continues_bit_identically(tmp_path):
    """{tables, grad rows, W0, epoch, LR} saved from the device mid-run; a NEW context that
    loads the file continues exactly as the uninterrupted run (deterministic mode: one block, file order)."""
    g = load_golden(TRAIN_CASES[0])
    N = g["clicks"].shape[0]

    def run(split):
        uw, iw, w0 = fresh_tables(g)
        ctx = make_ctx(g, deterministic=True, milestones=(1,))
        ctx.bind_tables(uw, iw, w0)
        ctx.train_epoch()                       # epoch 0, LR 1e-2
        ctx.train_range(0, N // 2)              # half of epoch 1: grad rows are non-zero here
        if split:
            path = str(tmp_path / "state.ckpt")
            ctx.save_state(path)
            ctx.close()
            uw, iw, w0 = (np.zeros_like(a) for a in (uw, iw, w0))
            ctx = make_ctx(g, deterministic=True, milestones=(1,))
            ctx.bind_tables(uw, iw, w0)
            ctx.load_state(path)
            assert ctx.get_epoch() == 1 and np.abs(uw).max() > 0       # bound host tables refreshed by the load
        ctx.train_range(N // 2, N)
        loss = ctx.train_epoch()                # epoch 1 again from the counter's point of view: LR schedule applies
        ctx.sync_to_host()
        gu, gi = ctx.grads()
        out = (uw.copy(), iw.copy(), w0.copy(), gu, gi, loss, ctx.get_lr(), ctx.get_epoch())
        ctx.close()
        return out

    a, b = run(False), run(True)
    for x, y in zip(a[:5], b[:5]):
        assert np.array_equal(x, y)
    assert a[5:] == b[5:]


def test_checkpoint_rejects_wrong_shape_and_garbage(tmp_path):
    g = load_golden(TRAIN_CASES[0])
    uw, iw, w0 = fresh_tables(g)
    ctx = make_ctx(g)
    ctx.bind_tables(uw, iw, w0)
    path = str(tmp_path / "s.ckpt")
    ctx.save_state(path)
    U, D = uw.shape
    other = abi.Context(D, int(g["K"]), U + 1, iw.shape[0])
    other.bind_tables(np.zeros((U + 1, D), np.float32), iw.copy(), w0.copy())
    with pytest.raises(abi.HeatError, match="shape differs"):
        other.load_state(path)
    bad = tmp_path / "bad.ckpt"
    bad.write_bytes(b"not a checkpoint" * 8)
    with pytest.raises(abi.HeatError, match="not a heat_b200 checkpoint"):
        ctx.load_state(str(bad))
    with open(path, "r+b") as f:
        f.truncate(64 + 100)
    with pytest.raises(abi.HeatError, match="truncated"):
        ctx.load_state(path)


TILE = dict(neg_sampler=1, tile_sampler=1, tile_size=32, refresh_interval=50)


@pytest.mark.parametrize("engine", [0, 1, 3])
def test_tile_sampler_indices_bit_exact_and_used_by_every_engine(engine):
    """SURVEY.md 8.6 row 2 (random_tile_negative_sampler.cpp:22-45 as a hot-tile policy): the device's tile-mode
    ids equal the oracle's restatement bit for bit, a refresh window only touches its tile, and each training kernel
    draws exactly those ids (three sequential steps match the oracle fed with them)."""
    g = load_golden("train_d64_k16")
    D, K, I = int(g["D"]), int(g["K"]), g["init_item_w"].shape[0]
    N = g["clicks"].shape[0]
    off = 777
    ctx = make_ctx(g, deterministic=True, w0_tail_mode=1, engine=engine, sample_offset=off, **TILE)
    dev = ctx.sample_negatives(2, 0, N)
    ref = oracle.neg_table(int(g["seed"]), 2, off, g["clicks"], K, I, tile_size=32, refresh_interval=50)
    assert np.array_equal(dev, ref)
    first = 50 - off % 50                                 # local index where the second refresh window starts
    window = ref[first:first + 50]
    assert len(set(window.ravel().tolist())) <= 32 + 2    # the tile (+ look-back values from the window before)
    assert not np.array_equal(ref, oracle.neg_table(int(g["seed"]), 2, off, g["clicks"], K, I))
    uw, iw, w0 = fresh_tables(g)
    ctx.bind_tables(uw, iw, w0)
    negs0 = ctx.sample_negatives(0, 0, 3)
    _, per = ctx.train_range(0, 3, per_sample=True)
    ctx.sync_to_host()
    a = fresh_tables(g)
    o = oracle.COracle(D, K, g["clicks"], g["his_items"], g["masks"], *a, clip_val=float(g["clip"]), l_r=float(g["lr"]))
    want = [o.step(int(g["clicks"][s, 0]), int(g["clicks"][s, 1]), negs0[s]) for s in range(3)]
    o.close()
    assert np.allclose(per, want, rtol=2e-3, atol=1e-5), (per, want)
    assert rel_err(uw, a[0]) < 2e-4 and rel_err(iw, a[1]) < 2e-4
    ctx.close()


def test_neg_sampler_1_without_tile_sampler_draws_uniformly():
    """The reference's shipped loop calls ignore_pos_sampling, which ignores the tile: neg_sampler 1
    alone must not change a single id."""
    g = load_golden("train_d64_k16")
    K, I, N = int(g["K"]), g["init_item_w"].shape[0], g["clicks"].shape[0]
    ctx = make_ctx(g, neg_sampler=1, tile_size=32, refresh_interval=50)
    assert np.array_equal(ctx.sample_negatives(0, 0, N), oracle.neg_table(int(g["seed"]), 0, 0, g["clicks"], K, I))


def test_auto_engine_choice():
    """engine 0: the warp-per-sample kernel (fresh-grad update, full grid) where it covers the shape, the large-negative
    kernel for big K, the CTA kernel for the sequential order and when asked for (engine 2 = the loss-strict choice)."""
    rng = np.random.default_rng(0)
    U, D, N, H = 64, 64, 4096, 100

    def engine_for(I, K=16, **kw):
        clicks = np.stack([np.sort(rng.integers(0, U, N)), rng.integers(0, I, N)], 1).astype(np.uint64)
        his = rng.integers(0, I, (U, H)).astype(np.uint64)
        masks = np.full((U, 1), H, np.uint64)
        ctx = abi.Context(D, K, U, I, **kw)
        ctx.set_dataset(clicks, his, masks)
        uw = (rng.standard_normal((U, D)) * 1e-2).astype(np.float32)
        iw = np.full((I, D), 1e-2, np.float32)
        ctx.bind_tables(uw, iw, (rng.standard_normal((D, D)) * 1e-2).astype(np.float32))
        assert ctx.last_engine() == 0
        loss = ctx.train_epoch()
        e = ctx.last_engine()
        ctx.close()
        assert np.isfinite(loss)
        return e

    assert engine_for(40000) == 1
    assert engine_for(300) == 1                            # a toy table: the default caps the warps in flight at one per 32 rows
    assert engine_for(40000, max_blocks=32) == 1
    assert engine_for(40000, engine=2) == 2
    assert engine_for(40000, deterministic=True) == 2
    assert engine_for(40000, K=48) == 2                    # outside the warp kernel's (emb_dim, num_negs): the CTA kernel covers it
    assert engine_for(40000, K=200) == 3


@pytest.mark.parametrize("update_mode", [0, 2, 3])
def test_update_modes_agree_when_nothing_races(update_mode):
    """One warp in flight (max_blocks 1, warps_per_block 1 is not offered, so: one block of 4 warps on disjoint users
    would still race on item rows) -- instead compare the Hogwild modes statistically on a small shape: same data, same
    negatives, two epochs; losses within 3 % of each other and of the sequential oracle's."""
    d = synth.generate_named("small")
    D, K = 64, 16
    a = synth.init_tables(d.num_users, d.num_items, D)
    ctx = abi.Context(D, K, d.num_users, d.num_items, engine=1, max_blocks=2, warps_per_block=4, update_mode=update_mode)
    ctx.set_dataset(d.clicks, d.his_items, d.masks); ctx.bind_tables(*a)
    got = [ctx.train_epoch(), ctx.train_epoch()]
    ctx.close()
    b = synth.init_tables(d.num_users, d.num_items, D)
    o = oracle.COracle(D, K, d.clicks, d.his_items, d.masks, *b)
    want = [o.train_epoch(nthreads=1), o.train_epoch(nthreads=1)]
    o.close()
    for g_, w_ in zip(got, want):
        assert abs(g_ - w_) / w_ < 0.05, (update_mode, got, want)


@pytest.mark.parametrize("engine,kw", [(0, {}), (1, dict(update_mode=3)), (2, {}), (2, dict(max_blocks=3)), (3, {})])
@pytest.mark.parametrize("lo,hi", [(0, 60000), (17, 18), (5, 70), (1000, 9001), (123, 58000)])
def test_every_sample_of_a_range_is_processed_exactly_once(engine, kw, lo, hi):
    """Guided scheduling (chunk_range: chunks of 64, then a tail in chunks of 8): whatever the range, every sample's loss
    slot is written once and the launch's loss is their sum -- on the Hogwild grid of all three kernel families."""
    d = synth.generate_named("small")
    D, K = 64, 16
    t = synth.init_tables(d.num_users, d.num_items, D)
    ctx = abi.Context(D, K, d.num_users, d.num_items, engine=engine, **kw)
    ctx.set_dataset(d.clicks, d.his_items, d.masks); ctx.bind_tables(*t)
    sentinel = np.float32(-7.0)
    out = np.full(hi - lo, sentinel, np.float32)
    import ctypes as C
    loss = C.c_double()
    abi._ck(abi.lib().heat_train_range(ctx._h, lo, hi, C.byref(loss), out.ctypes.data_as(C.POINTER(C.c_float))))
    ctx.close()
    assert (out > 0).all() and np.isfinite(out).all(), int((out == sentinel).sum())
    assert abs(float(out.astype(np.float64).sum()) - loss.value) <= 1e-3 * max(1.0, abs(loss.value))
