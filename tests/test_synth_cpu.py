"""Host-side generators behind the benchmarks and the multi-rank tests (no GPU): the weak-scaling form of a shape, the
user blocks of the structured power-law dataset of BASELINE configs[4], and their union."""
import numpy as np

from heat_b200 import shard, synth


def _check_clickdata(d):
    assert d.clicks.dtype == np.uint64 and d.clicks.shape[1] == 2
    assert d.his_items.shape == (d.num_users, d.max_his) and d.masks.shape == (d.num_users, 1)
    assert int(d.clicks[:, 0].max()) < d.num_users and int(d.clicks[:, 1].max()) < d.num_items
    assert d.train_indptr[-1] == d.train_size == d.train_items.size
    assert np.all(np.diff(d.clicks[:, 0].astype(np.int64)) >= 0)            # a user's samples are consecutive
    assert np.array_equal(np.bincount(d.clicks[:, 0].astype(np.int64), minlength=d.num_users), np.diff(d.train_indptr))
    assert int(d.masks.max()) <= d.max_his and int(d.his_items.max()) < d.num_items


def test_weak_scaling_form_shares_the_item_catalogue():
    a = synth.generate_named("tiny")
    b = synth.generate_named("tiny", users_mult=3)
    _check_clickdata(a); _check_clickdata(b)
    assert b.num_users == 3 * a.num_users and b.num_items == a.num_items and b.train_size == 3 * a.train_size
    # user-range shards of the union tile it, with the global sample offsets the Philox counter needs
    off = 0
    for r in range(3):
        sh, o = shard.shard_clicks(b, r, 3)
        _check_clickdata(sh)
        assert o == off
        off += sh.train_size
    assert off == b.train_size


def test_powerlaw_blocks_are_one_dataset():
    kw = dict(users_per_block=3000, num_items=5000, samples_per_block=60000, seed=7)
    b0, b1 = synth.generate_powerlaw_block(0, **kw), synth.generate_powerlaw_block(1, **kw)
    again = synth.generate_powerlaw_block(1, **kw)
    _check_clickdata(b0); _check_clickdata(b1)
    assert np.array_equal(b1.clicks, again.clicks) and np.array_equal(b1.his_items, again.his_items)     # a block depends on (seed, block) only
    assert not np.array_equal(b0.clicks[:1000], b1.clicks[:1000])
    for b in (b0, b1):          # evaluation masks: every user's train items sorted, disjoint from the test items
        for u in range(0, b.num_users, 97):
            tr = b.train_items[b.train_indptr[u]:b.train_indptr[u + 1]]
            te = b.test_items[b.test_indptr[u]:b.test_indptr[u + 1]]
            assert np.all(np.diff(tr) > 0) and not np.intersect1d(tr, te).size
    # shared item side: the blocks' item popularity is strongly correlated (same popularity law and clusters)
    p0 = np.bincount(b0.train_items, minlength=5000).astype(float); p1 = np.bincount(b1.train_items, minlength=5000).astype(float)
    assert np.corrcoef(p0, p1)[0, 1] > 0.8
    un = synth.concat_blocks([b0, b1])
    _check_clickdata(un)
    assert un.num_users == 6000 and un.train_size == b0.train_size + b1.train_size
    assert np.array_equal(un.clicks[b0.train_size:, 1], b1.clicks[:, 1])
    assert int(un.clicks[b0.train_size, 0]) == 3000 + int(b1.clicks[0, 0])
