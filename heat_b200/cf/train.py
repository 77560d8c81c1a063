"""Engine wrapper (reference: cf_cpu/cf/train.py:4-24) plus the fused GPU evaluation."""
import numpy as np

try:
    from .cpp_base import CPPBase
    from . import cf_c
    from . import metrics as host_metrics
except ImportError:
    from cpp_base import CPPBase
    import cf_c
    import metrics as host_metrics

METRIC_ORDER = ["Recall(k=20)", "Recall(k=50)", "NDCG(k=20)", "NDCG(k=50)", "HitRate(k=20)", "HitRate(k=50)"]


class Engine(CPPBase):
    def __init__(self, dataset=None, aggregator_weights=None, model=None, cf_config=None):
        super().__init__()
        self.c_class = cf_c.modules.train.Engine
        self._keep = (dataset, aggregator_weights, model, cf_config)   # cf_c borrows their buffers
        self.init_c_instance(dataset=dataset.c_instance, aggregator_weights=aggregator_weights.c_instance,
                             model=model.c_instance, cf_config=cf_config.c_instance)

    def train_one_epoch(self):
        return self.c_instance.train_one_epoch()

    def save_checkpoint(self, path):
        """{user/item tables, persistent grad rows, W0, epoch, scheduled LR} -> one file, read from the device."""
        self.c_instance.save_checkpoint(str(path))

    def load_checkpoint(self, path):
        """Restore a save_checkpoint file: the device state and the numpy tables the model objects alias."""
        self.c_instance.load_checkpoint(str(path))

    def evaluate0(self):
        """Dense [num_users, num_items] fp32 score matrix (compatibility path, engine.cpp:388-400)."""
        return self.c_instance.evaluate0()

    def evaluate_metrics(self, train_data, test_data, metrics=None, reduce_over_ranks=False):
        """Score GEMM + masked top-50 + Recall/NDCG/HitRate@20/50 on the GPU; the score matrix is
        never materialised.  Returns {metric: value} for the requested names, printed in the
        reference's `[Metrics] ...` format (metrics.py:35)."""
        metrics = [m.replace(" ", "") for m in metrics] if metrics else METRIC_ORDER
        tr_ip, tr_it = train_data.csr()
        te_ip, te_it = test_data.csr(num_users=train_data.num_users)
        if any(m not in METRIC_ORDER for m in metrics):
            # the rest of the reference's list (Precision, F1, MRR, MAP, NormalizedRecall, DCG, other cut-offs,
            # metrics.py:51-158): fused GEMM + top-k on the GPU, the per-user formulas on the host from the id lists
            if reduce_over_ranks:
                raise NotImplementedError("metrics outside %s are not reduced over ranks" % METRIC_ORDER)
            max_k = max(host_metrics.parse_metric(m)[1] for m in metrics)
            ids, _ = self.c_instance.evaluate_topk(max_k, tr_ip, tr_it)
            res = host_metrics.metrics_from_topk(ids, te_ip, te_it, metrics)
            print("[Metrics] " + " - ".join("{}: {:.6f}".format(k, v) for k, v in res.items()))
            return res
        out, sums = self.c_instance.evaluate_metrics(tr_ip, tr_it, te_ip, te_it)
        if reduce_over_ranks:
            sums = self.c_instance.allreduce(sums)
            out = sums[:6] / max(sums[6], 1.0)
        res = dict(zip(METRIC_ORDER, np.asarray(out).tolist()))
        res = {m: res[m] for m in metrics}
        print("[Metrics] " + " - ".join("{}: {:.6f}".format(k, v) for k, v in res.items()))
        return res
