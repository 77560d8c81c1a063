"""heat_train_range_noalloc (static device scratch, no cudaMalloc) against heat_train_range, through the C ABI.

Collected last on purpose.  Passes alone in a fresh process and after the other tests:
it only depends on the golden fixture and on the library."""
import numpy as np
import pytest

from heat_b200 import abi
from helpers import fresh_tables, load_golden, rel_err

pytestmark = pytest.mark.gpu


def _ctx(g, **kw):
    U, D = g["init_user_w"].shape
    I = g["init_item_w"].shape[0]
    ctx = abi.Context(D, int(g["K"]), U, I, clip_val=float(g["clip"]), l_r=float(g["lr"]), seed=int(g["seed"]), **kw)
    ctx.set_dataset(g["clicks"], g["his_items"], g["masks"])
    return ctx


@pytest.mark.parametrize("engine", [0, 1, 3])
def test_noalloc_matches_train_range_two_calls_in_a_row(engine):
    g = load_golden("train_d64_k16")
    N = g["clicks"].shape[0]
    half = N // 2
    a_t, b_t = fresh_tables(g), fresh_tables(g)
    a = _ctx(g, deterministic=True, w0_tail_mode=1, engine=engine)
    b = _ctx(g, deterministic=True, w0_tail_mode=1, engine=engine)
    a.bind_tables(*a_t)
    b.bind_tables(*b_t)
    # two calls in a row on the shared static scratch: the second must not see the first one's counter / sum / losses
    for lo, hi in ((0, half), (half, N)):
        want_sum, want = a.train_range(lo, hi, per_sample=True)
        got_sum, got = b.train_range_noalloc(lo, hi, per_sample=True)
        assert np.allclose(got, want, rtol=1e-5, atol=1e-6), np.abs(got - want).max()
        assert abs(got_sum - want_sum) <= 1e-5 * max(1.0, abs(want_sum))
        assert a.last_engine() == b.last_engine()
    # without the per-sample output only the 16-byte header of the scratch is used
    want_sum = a.train_range(0, 64)
    got_sum = b.train_range_noalloc(0, 64)
    assert abs(got_sum - want_sum) <= 1e-5 * max(1.0, abs(want_sum))
    a.sync_to_host(); b.sync_to_host()
    for x, y in zip(a_t, b_t):
        assert rel_err(y, x) < 1e-5
    ga, gb = a.grads(), b.grads()
    for x, y in zip(ga, gb):
        assert rel_err(y, x) < 1e-4
    a.close(); b.close()


def test_noalloc_hogwild_grid_and_empty_range():
    g = load_golden("train_d64_k16")
    N = g["clicks"].shape[0]
    t = fresh_tables(g)
    c = _ctx(g, max_blocks=4)
    c.bind_tables(*t)
    assert c.train_range_noalloc(5, 5) == 0.0
    l0, per0 = c.train_range_noalloc(0, N, per_sample=True)
    c.zero_grad()
    l1, per1 = c.train_range_noalloc(0, N, per_sample=True)
    assert np.isfinite(per0).all() and np.isfinite(per1).all()
    assert abs(float(per0.astype(np.float64).sum()) - l0) < 1e-3 * max(1.0, abs(l0))
    assert abs(float(per1.astype(np.float64).sum()) - l1) < 1e-3 * max(1.0, abs(l1))
    assert l1 < l0
    c.close()


def test_noalloc_rejects_ranges_beyond_the_static_scratch():
    U, I, D, K, H = 64, 128, 64, 16, 4
    N = abi.NOALLOC_MAX_SAMPLES + 1
    rng = np.random.default_rng(3)
    clicks = np.stack([np.sort(rng.integers(0, U, N)), rng.integers(0, I, N)], 1).astype(np.uint64)
    his = rng.integers(0, I, (U, H)).astype(np.uint64)
    masks = np.full((U, 1), H, np.uint64)
    c = abi.Context(D, K, U, I, max_blocks=8)
    c.set_dataset(clicks, his, masks)
    c.bind_tables((rng.standard_normal((U, D)) * 1e-2).astype(np.float32), (rng.standard_normal((I, D)) * 1e-2).astype(np.float32),
                  (rng.standard_normal((D, D)) * 1e-2).astype(np.float32))
    with pytest.raises(abi.HeatError, match="HEAT_NOALLOC_MAX_SAMPLES"):
        c.train_range_noalloc(0, N)
    with pytest.raises(abi.HeatError, match="out of bounds"):
        c.train_range_noalloc(0, N + 1)
    assert np.isfinite(c.train_range_noalloc(1, N))        # exactly the maximum is accepted
    c.close()
