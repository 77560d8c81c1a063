"""CPU tests of the `cf_c` module surface and of the Python frontend (heat_b200/cf): same names and
keywords as the reference binding (init_modules.cpp:13-152), parsing / history / sharding, and the
host metrics against golden values from the reference's cf/metrics.py."""
import os

import numpy as np
import pytest

from heat_b200 import synth
from heat_b200.cf import cf_c, cf_config, datasets, metrics
from helpers import load_golden


def test_cf_c_surface_matches_reference_binding():
    m = cf_c.modules
    c = m.CFConfig(emb_dim=64, num_negs=16, num_users=10, num_items=20, train_size=5, neg_sampler=0, tile_size=512,
                   refresh_interval=8192, num_subepoches=2, l2=1e-7, clip_val=1.0, milestones=[10], l_r=0.01)
    assert c.emb_dim == 64
    c.emb_dim = 32; assert c.emb_dim == 32; c.emb_dim = 64
    ds = m.datasets.ClickDataset(click_dataset=np.zeros((5, 2), np.uint64), historical_items=np.zeros((10, 4), np.uint64),
                                 masks=np.ones((10, 1), np.uint64))
    assert isinstance(ds, m.datasets.Dataset) and ds.data_rows == 5
    ds.max_his = 4; assert ds.max_his == 4
    mf = m.models.MatrixFactorization(cf_config=c, user_weights=np.zeros((10, 64), np.float32), item_weights=np.zeros((20, 64), np.float32))
    assert isinstance(mf, m.models.Model)
    aw = m.behavior_aggregators.AggregatorWeights(aggregator_weights0=np.zeros((64, 64), np.float32))
    assert aw.emb_dim == 64
    # additive GPU-path knobs: plain attributes with the library defaults (0), copied by cf_config.CFConfig.init_c_instance
    for knob in cf_config.GPU_KNOBS:
        assert hasattr(c, knob), knob
        if knob != "seed":
            assert getattr(c, knob) == 0, knob
    py = cf_config.CFConfig(emb_dim=64, num_negs=16, engine=2, sync_average=2, sync_grads=1, eval_precision=1, w0_tail_mode=1)
    py.num_users, py.num_items, py.train_size = 10, 20, 5
    py.init_c_instance()
    assert (py.c_instance.engine, py.c_instance.sync_average, py.c_instance.sync_grads, py.c_instance.eval_precision,
            py.c_instance.w0_tail_mode) == (2, 2, 1, 1, 1)
    with pytest.raises(TypeError):
        cf_config.CFConfig(emb_dim=64, no_such_knob=1)
    m.test.test_out().test("cf_c surface ok")
    for name in ("train_one_epoch", "evaluate0", "evaluate_topk", "evaluate_metrics", "sync_to_host", "comm_init"):
        assert hasattr(m.train.Engine, name)
    with pytest.raises(Exception):
        m.models.MatrixFactorization(cf_config=c, user_weights=np.zeros((9, 64), np.float32), item_weights=np.zeros((20, 64), np.float32))
    import torch
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CPU fallback"):
            m.train.Engine(dataset=ds, aggregator_weights=aw, model=mf, cf_config=c)


def test_frontend_parses_lightgcn_files_like_the_generator(tmp_path):
    d = synth.generate(300, 500, 9000, max_his=20, seed=4)
    synth.write_lightgcn_txt(d, str(tmp_path / "train.txt"), str(tmp_path / "test.txt"))
    cfg = cf_config.CFConfig(emb_dim=32, num_negs=8, max_his=20, seed=4)
    tr = datasets.ClickDataset(str(tmp_path / "train.txt"), config=cfg)
    te = datasets.ClickDataset(str(tmp_path / "test.txt"), config=cfg)
    assert (cfg.num_users, cfg.train_size) == (300, 9000) and cfg.num_items >= int(d.train_items.max()) + 1
    assert np.array_equal(tr.click_dataset, d.clicks) and tr.click_dataset.dtype == np.uint64
    assert np.array_equal(tr.masks, d.masks) and np.array_equal(tr.his_items, d.his_items)
    assert tr.c_instance.data_rows == 9000 and tr.c_instance.max_his == 20
    assert te.c_instance is None                       # only the train set reaches cf_c (datasets.py:70-79)
    ip, it = te.csr(num_users=300)
    assert np.array_equal(ip, d.test_indptr) and np.array_equal(it, d.test_items)
    sub = datasets.SubClickDataset(tr, 100, 200, 1)
    assert sub.num_users == 100 and sub.sample_offset == int(d.train_indptr[100])
    assert np.array_equal(sub.click_dataset[:, 1], d.clicks[d.train_indptr[100]:d.train_indptr[200], 1])
    assert sub.click_dataset[:, 0].min() == 0 and sub.click_dataset[:, 0].max() == 99
    assert np.array_equal(sub.his_items, d.his_items[100:200])


def test_host_metrics_match_reference_metrics_py_golden():
    g = load_golden("eval_d64")

    class DS:
        def __init__(self, indptr, items):
            self.indptr, self.items = indptr, items

        def csr(self, num_users=None):
            return self.indptr, self.items
    res = metrics.evaluate_metrics(DS(g["train_indptr"], g["train_items"]), DS(g["test_indptr"], g["test_items"]),
                                   g["scores"].copy(), [str(n) for n in g["metric_names"]])
    for n, v in zip(g["metric_names"], g["metric_values"]):
        assert abs(res[str(n)] - float(v)) < 1e-9, (n, res[str(n)], float(v))
    with pytest.raises(NotImplementedError):
        metrics.parse_metric("AUC(k=3)")


@pytest.mark.skipif(not os.path.isdir("/root/reference/cf_cpu/cf"), reason="reference not present")
def test_all_nine_host_metrics_match_the_reference_module():
    import contextlib, importlib.util, io
    spec = importlib.util.spec_from_file_location("refmetrics", "/root/reference/cf_cpu/cf/metrics.py")
    rm = importlib.util.module_from_spec(spec); spec.loader.exec_module(rm)
    d = synth.generate(200, 400, 6000, max_his=20, seed=8)
    sim = np.random.default_rng(0).standard_normal((200, 400)).astype(np.float32)
    names = ['Recall(k=20)', 'NDCG(k=50)', 'HitRate(k=20)', 'Precision(k=10)', 'F1(k=20)', 'MRR(k=50)', 'MAP(k=50)',
             'NormalizedRecall(k=20)', 'DCG(k=20)']

    class DS:
        def __init__(self, indptr, items):
            self.indptr, self.items = indptr, items
            self.user_items_dic = {u: items[indptr[u]:indptr[u + 1]].tolist() for u in range(200) if indptr[u + 1] > indptr[u]}

        def csr(self, num_users=None):
            return self.indptr, self.items
    a, b = DS(d.train_indptr, d.train_items), DS(d.test_indptr, d.test_items)
    mine = metrics.evaluate_metrics(a, b, sim.copy(), names)
    with contextlib.redirect_stdout(io.StringIO()):
        ref = rm.evaluate_metrics(a, b, sim.copy(), names)
    for n in names:
        assert abs(mine[n] - ref[n]) < 1e-9, (n, mine[n], ref[n])


def test_native_click_parser_matches_the_reference_line_reader(tmp_path):
    """heat_clickfile_* (C ABI, one thread per core) against the reference's per-line python
    reader (datasets.py:31-40) -- blank lines, tabs / CRLF, a custom separator, unsorted users, a user without
    items, a file large enough to be cut into per-thread chunks, and the error path."""
    from heat_b200 import abi
    from heat_b200.cf import datasets
    rng = np.random.default_rng(5)
    small = tmp_path / "small.txt"
    small.write_text("3 10 11 12\n\n0 5\n2\n1 7 8\t9 \r\n")
    for got in (abi.parse_clickfile(str(small)), datasets.cf_c.modules.datasets.parse_clicks(str(small), " ")):
        want = datasets.parse_python(str(small))
        for g, w in zip(got, want):
            assert np.array_equal(g, w)
    assert abi.parse_clickfile(str(small))[0].tolist() == [3, 0, 2, 1]
    comma = tmp_path / "comma.txt"
    comma.write_text("0,4,5\n1,6\n")
    u, ip, it = abi.parse_clickfile(str(comma), ",")
    assert u.tolist() == [0, 1] and ip.tolist() == [0, 2, 3] and it.tolist() == [4, 5, 6]
    big = tmp_path / "big.txt"
    with open(big, "w") as f:
        for user in range(60000):
            f.write(" ".join([str(user)] + [str(int(x)) for x in rng.integers(0, 10**6, int(rng.integers(1, 40)))]) + "\n")
    assert big.stat().st_size > (1 << 20)
    for g, w in zip(abi.parse_clickfile(str(big)), datasets.parse_python(str(big))):
        assert np.array_equal(g, w)
    bad = tmp_path / "bad.txt"
    bad.write_text("0 1 2\n1 x 3\n")
    with pytest.raises(abi.HeatError, match=r"bad.txt:2: unexpected character 'x'"):
        abi.parse_clickfile(str(bad))
    with pytest.raises(abi.HeatError, match="cannot open"):
        abi.parse_clickfile(str(tmp_path / "missing.txt"))


def test_history_arrays_follow_the_reference_rule():
    """datasets.py:47-61: >= max_his items -> a sample of max_his of them (no repeats, all from the user's list);
    fewer -> all items in order padded with the last; mask = min(len, max_his)."""
    from heat_b200.cf import datasets
    rng = np.random.default_rng(1)
    n = np.array([0, 1, 5, 8, 9, 30], dtype=np.int64)
    indptr = np.concatenate([[0], np.cumsum(n)])
    items = rng.permutation(1000)[:indptr[-1]].astype(np.int32)
    his, masks = datasets._history(indptr, items, 8, seed=3)
    assert masks[:, 0].tolist() == [0, 1, 5, 8, 8, 8]
    assert his[1].tolist() == [items[0]] * 8
    assert his[2].tolist() == items[1:6].tolist() + [items[5]] * 3
    for u in (3, 4, 5):
        row = items[indptr[u]:indptr[u + 1]]
        assert len(set(his[u].tolist())) == 8 and set(his[u].tolist()) <= set(row.tolist())
    again, _ = datasets._history(indptr, items, 8, seed=3)
    assert np.array_equal(his, again)


def test_sub_click_datasets_tile_the_parent(tmp_path):
    """The multi-rank frontend: user-range shards (main.py:51-66) cover every sample once, keep item ids global, re-base user
    ids, and carry the global index of their first sample (the Philox counter base)."""
    from heat_b200.cf import main as cf_main
    d = synth.generate(211, 300, 6000, max_his=20, seed=9)
    synth.write_lightgcn_txt(d, str(tmp_path / "train.txt"), str(tmp_path / "test.txt"))
    cfg = cf_config.CFConfig(emb_dim=32, num_negs=8, max_his=20)
    parent = datasets.ClickDataset(str(tmp_path / "train.txt"), config=cfg)
    world, seen, offset = 3, 0, 0
    for r in range(world):
        a, b = cf_main.user_range(parent.num_users, r, world)
        sub = datasets.SubClickDataset(parent, a, b, r)
        assert sub.num_users == b - a and sub.num_items == parent.num_items
        assert sub.sample_offset == offset
        n = sub.click_dataset.shape[0]
        assert np.array_equal(sub.click_dataset[:, 1], parent.click_dataset[offset:offset + n, 1])
        assert np.array_equal(sub.click_dataset[:, 0] + np.uint64(a), parent.click_dataset[offset:offset + n, 0])
        assert np.array_equal(sub.his_items, parent.his_items[a:b]) and np.array_equal(sub.masks, parent.masks[a:b])
        ip, it = sub.csr()
        assert ip[-1] == n and np.array_equal(it, parent.items[parent.indptr[a]:parent.indptr[b]])
        seen += n; offset += n
    assert seen == parent.click_dataset.shape[0]
