"""Host (numpy) ranking metrics with the reference's definitions (cf_cpu/cf/metrics.py:5-159).

`evaluate_metrics(train_data, test_data, sim_matrix, metrics)` keeps the reference's signature and
`[Metrics] ...` output line for callers that still hold a dense score matrix (Engine.evaluate0()).
The training driver does not use it: Engine.evaluate_metrics runs GEMM + top-k + Recall/NDCG/
HitRate on the GPU.  `metrics_from_topk` computes every metric the reference defines from top-k id
lists, e.g. the ones Engine.evaluate_topk returns.

Definitions (k = cut-off, T = the user's test items, hits = ranks r < k whose item is in T):
  Recall |hits|/(|T|+1e-12) :44-48      NormalizedRecall |hits|/min(k,|T|+1e-12) :57-61
  Precision |hits|/(k+1e-12) :70-74     F1 2PR/(P+R+1e-12) :80-84
  DCG sum 1/ln(2+r) :93-100             NDCG DCG/(DCG of the first min(k,|T|) ranks + 1e-12) :108-113
  MRR sum 1/(r+1) (sic: summed over all hits) :122-128      HitRate [|hits|>0] :135-139
  MAP (sum over hits of (#hits so far)/(r+1)) / (|hits|+1e-12) :148-158
"""
import re

import numpy as np

_NAME = re.compile(r"^(Recall|NormalizedRecall|Precision|F1|DCG|NDCG|MRR|HitRate|MAP)\(k=(\d+)\)$")


def parse_metric(name):
    m = _NAME.match(name.replace(" ", ""))
    if not m:
        raise NotImplementedError('metrics={} not implemented.'.format(name))
    return m.group(1), int(m.group(2))


def _one(kind, k, hit, n_true):
    """hit: bool [max_k] (is the item at rank r a test item), n_true = |T|."""
    h = hit[:k]
    nh = int(h.sum())
    ranks = np.nonzero(h)[0]
    if kind == "Recall":
        return nh / (n_true + 1e-12)
    if kind == "NormalizedRecall":
        return nh / min(k, n_true + 1e-12)
    if kind == "Precision":
        return nh / (k + 1e-12)
    if kind == "F1":
        p, r = nh / (k + 1e-12), nh / (n_true + 1e-12)
        return 2 * p * r / (p + r + 1e-12)
    if kind == "DCG":
        return float(np.sum(1.0 / np.log(2.0 + ranks)))
    if kind == "NDCG":
        idcg = float(np.sum(1.0 / np.log(2.0 + np.arange(min(k, n_true)))))
        return float(np.sum(1.0 / np.log(2.0 + ranks))) / (idcg + 1e-12)
    if kind == "MRR":
        return float(np.sum(1.0 / (ranks + 1.0)))
    if kind == "HitRate":
        return 1 if nh > 0 else 0
    if kind == "MAP":
        return float(np.sum(np.arange(1, nh + 1) / (ranks + 1.0))) / (nh + 1e-12)
    raise NotImplementedError(kind)


def metrics_from_topk(top_k_items, test_indptr, test_items, metrics):
    """Mean of each metric over the users that have test items.  top_k_items: int [U, max_k], sorted by
    score descending (ids < 0 = empty slot)."""
    parsed = [parse_metric(m) for m in metrics]
    max_k = max(k for _, k in parsed)
    if top_k_items.shape[1] < max_k:
        raise ValueError(f"need top-{max_k} lists, got top-{top_k_items.shape[1]}")
    sums, n = np.zeros(len(parsed)), 0
    for u in range(len(test_indptr) - 1):
        true = test_items[test_indptr[u]:test_indptr[u + 1]]
        if true.size == 0:
            continue
        n += 1
        hit = np.isin(top_k_items[u, :max_k], true) & (top_k_items[u, :max_k] >= 0)
        for j, (kind, k) in enumerate(parsed):
            sums[j] += _one(kind, k, hit, true.size)      # the list length, duplicates included (metrics.py:57-61,108-113; metrics_kernel)
    return dict(zip(metrics, (sums / max(n, 1)).tolist()))


def evaluate_metrics(train_data, test_data, sim_matrix, metrics):
    """Reference signature (metrics.py:5): masks the train items of every test user to -inf IN PLACE,
    takes the top max_k per row, averages the metrics over the test users, prints and returns them."""
    print(f'Evaluating metrics {metrics} ...')
    max_k = max(parse_metric(m)[1] for m in metrics)
    tr_ip, tr_it = train_data.csr()
    te_ip, te_it = test_data.csr(num_users=sim_matrix.shape[0])
    for u in range(sim_matrix.shape[0]):
        if te_ip[u + 1] > te_ip[u]:
            sim_matrix[u, tr_it[tr_ip[u]:tr_ip[u + 1]]] = -np.inf
    part = np.argpartition(-sim_matrix, max_k - 1, axis=1)[:, :max_k]
    order = np.argsort(-np.take_along_axis(sim_matrix, part, axis=1), axis=1, kind="stable")
    top = np.take_along_axis(part, order, axis=1)
    res = metrics_from_topk(top, te_ip, te_it, metrics)
    print('[Metrics] ' + ' - '.join('{}: {:.6f}'.format(k, v) for k, v in res.items()))
    return res
