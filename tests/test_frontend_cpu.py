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
