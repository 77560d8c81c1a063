"""GPU end-to-end test of the drop-in boundary: the Python frontend (same classes / CLI / YAML as the
reference's cf_cpu/cf) driving the `cf_c` module on LightGCN-format files."""
import numpy as np
import pytest
import yaml

from heat_b200 import synth

pytestmark = pytest.mark.gpu


def write_case(tmp_path, epochs=5, **gpu):
    d = synth.generate_named("small", max_his=100, seed=2022)
    synth.write_lightgcn_txt(d, str(tmp_path / "train.txt"), str(tmp_path / "test.txt"))
    cfg = {"dataset_config": {"data_dir": str(tmp_path), "data_format": "txt", "separator": " ",
                              "train_data": "train.txt", "test_data": "test.txt"},
           "model_config": {"num_negs": 16, "max_his": 100, "embedding_dim": 64, "neg_sampler": 0, "tile_size": 512,
                            "refresh_interval": 8192, "embedding_dropout": 0, "similarity_score": "dot",
                            "loss": "PairwiseLogisticLoss",
                            "metrics": ['Recall(k=20)', 'Recall(k=50)', 'NDCG(k=20)', 'NDCG(k=50)', 'HitRate(k=20)', 'HitRate(k=50)'],
                            "optimizer": "sgd", "milestones": [10], "learning_rate": 1.e-2, "clip_val": 1.0,
                            "embedding_regularizer": 1.e-7, "net_regularizer": 0, "net_dropout": 0, "epochs": epochs,
                            "eval_interval": 2, "seed": 2022}}
    if gpu:
        cfg["gpu"] = gpu
    p = tmp_path / "config0.yaml"
    p.write_text(yaml.safe_dump(cfg))
    return d, str(p)


def test_main_trains_and_evaluates_from_yaml(tmp_path, capsys):
    from heat_b200.cf import main as cf_main
    d, cfg = write_case(tmp_path, max_blocks=8, warps_per_block=4)
    res = cf_main.main(["--config", cfg])
    out = capsys.readouterr().out
    assert "[Metrics] Recall(k=20):" in out and "epoch: 4; loss:" in out
    assert set(res) == {'Recall(k=20)', 'Recall(k=50)', 'NDCG(k=20)', 'NDCG(k=50)', 'HitRate(k=20)', 'HitRate(k=50)'}
    assert res['Recall(k=20)'] > 0.05 and res['Recall(k=50)'] >= res['Recall(k=20)']
    dense = cf_main.main(["--config", cfg, "--dense-eval"])      # the reference's route: evaluate0 + numpy metrics
    # different torch-seeded runs of a Hogwild trainer: same quality, not the same bits
    assert abs(dense['Recall(k=20)'] - res['Recall(k=20)']) < 0.03


def test_in_place_numpy_contract_and_eval_paths_agree(tmp_path):
    """after train_one_epoch() the arrays the frontend handed over ARE the trained tables;
    evaluate0() + host metrics and the fused GPU metrics agree on them."""
    from heat_b200.cf import cf_config, datasets, metrics
    from heat_b200.cf.behavior_aggregators import AggregatorWeights
    from heat_b200.cf.models import MatrixFactorization
    from heat_b200.cf.train import Engine
    d, _ = write_case(tmp_path)
    cfg = cf_config.CFConfig(emb_dim=64, num_negs=16, max_his=100, neg_sampler=0, tile_size=512, refresh_interval=8192,
                             l2=1e-7, clip_val=1.0, milestones=[10], l_r=1e-2, max_blocks=8, warps_per_block=4)
    tr = datasets.ClickDataset(str(tmp_path / "train.txt"), config=cfg)
    te = datasets.ClickDataset(str(tmp_path / "test.txt"), config=cfg)
    cfg.init_c_instance()
    aw, model = AggregatorWeights(cfg), MatrixFactorization(cfg)
    model.init_c_instance(cfg)
    before = model.user_weights.copy()
    eng = Engine(tr, aw, model, cfg)
    losses = [eng.train_one_epoch() for _ in range(3)]
    assert losses[2] < losses[0] and not np.array_equal(before, model.user_weights)
    assert np.shares_memory(model.user_weights, model.user_embedding.weight.detach().numpy())
    names = ['Recall(k=20)', 'NDCG(k=20)', 'HitRate(k=50)']
    fused = eng.evaluate_metrics(tr, te, names)
    sim = eng.evaluate0()
    assert sim.shape == (cfg.num_users, cfg.num_items)
    ref = model.user_weights.astype(np.float64) @ model.item_weights.astype(np.float64).T
    assert np.abs(sim - ref).max() / np.abs(ref).max() < 1e-5
    host = metrics.evaluate_metrics(tr, te, sim, names)
    for n in names:
        assert abs(fused[n] - host[n]) < 1e-5, (n, fused[n], host[n])
    ids, sc = eng.c_instance.evaluate_topk(50, *tr.csr())
    allm = metrics.metrics_from_topk(ids, *te.csr(num_users=cfg.num_users), ['MRR(k=50)', 'MAP(k=50)', 'Precision(k=10)'])
    assert all(np.isfinite(v) for v in allm.values())


def test_cli_checkpoint_resume_and_full_metric_list(tmp_path, capsys):
    """--checkpoint / --resume and the YAML `metrics` list beyond Recall/NDCG/HitRate
    (row 4: Precision, F1, MRR, MAP, NormalizedRecall from the fused top-k lists)."""
    from heat_b200.cf import main as cf_main
    d, cfg = write_case(tmp_path, epochs=3, max_blocks=8, warps_per_block=4)
    ck = str(tmp_path / "run.ckpt")
    cf_main.main(["--config", cfg, "--checkpoint", ck])
    capsys.readouterr()
    import yaml as _y
    c = _y.safe_load(open(cfg))
    c["model_config"]["epochs"] = 5
    c["model_config"]["metrics"] = ['Recall(k=20)', 'Precision(k=10)', 'F1(k=20)', 'MRR(k=50)', 'MAP(k=50)', 'NormalizedRecall(k=20)']
    open(cfg, "w").write(_y.safe_dump(c))
    res = cf_main.main(["--config", cfg, "--resume", ck])
    out = capsys.readouterr().out
    assert "Resumed from" in out and "at epoch 3" in out and "epoch: 3; loss:" in out and "epoch: 0; loss:" not in out
    assert set(res) == set(c["model_config"]["metrics"])
    assert res['Recall(k=20)'] > 0.05 and all(np.isfinite(v) for v in res.values())
