"""Seeded synthetic click datasets at the BASELINE.json shapes.

The reference trains on LightGCN-format files (``user item item ...`` per line,
cf_cpu/cf/datasets.py:31-40) that are not available offline, so every test and
benchmark uses data of the same SHAPE generated here: power-law item popularity,
log-normal user degrees, and a planted cluster structure so that Recall@k is
non-trivial.  The history/mask arrays follow cf_cpu/cf/datasets.py:47-61 exactly
(``random.sample`` when a user has >= max_his items, else pad with the last item).

Nothing here is on the timed path; it only produces the host arrays that the
reference frontend would hand to ``cf_c`` (uint64 clicks [N,2], his_items
[U,max_his], masks [U,1]; float32 tables initialised N(0, 1e-2), models.py:15-16).
"""
from __future__ import annotations

import dataclasses
import random as _pyrandom

import numpy as np

# (users, items, train interactions) -- BASELINE.json configs / SURVEY.md 8.0
SHAPES = {
    "gowalla": (29858, 40981, 810128),
    "yelp18": (31668, 38048, 1237259),
    "amazon-book": (52643, 91599, 2380730),
    # large catalogues (sparser than the LightGCN sets): the parity budget of one sample in flight per ~600 item rows
    # reaches a grid of the warp-per-sample kernel here (train_launch.cu: auto_warp_engine) -- 8 resp. 12 warps per SM
    "wide-750k": (100000, 750000, 5000000),
    "wide-1100k": (100000, 1100000, 5000000),
    "tiny": (200, 300, 4000),
    "small": (2000, 3000, 60000),
}


@dataclasses.dataclass
class ClickData:
    num_users: int
    num_items: int
    max_his: int
    clicks: np.ndarray        # [N, 2] uint64, file order (all of a user's samples consecutive)
    his_items: np.ndarray     # [U, max_his] uint64
    masks: np.ndarray         # [U, 1] uint64
    train_indptr: np.ndarray  # [U+1] int64  CSR of train items per user (for eval masking)
    train_items: np.ndarray   # [N] int32
    test_indptr: np.ndarray   # [U+1] int64
    test_items: np.ndarray    # [T] int32

    @property
    def train_size(self) -> int:
        return int(self.clicks.shape[0])


def _degrees(rng, num_users, total, cap):
    if num_users * cap < total or total < num_users:
        raise ValueError(f"cannot place {total} interactions on {num_users} users with max degree {cap}")
    raw = rng.lognormal(mean=0.0, sigma=1.0, size=num_users)
    deg = np.maximum(1, np.floor(raw * (total / raw.sum())).astype(np.int64))
    deg = np.minimum(deg, cap)
    # fix the sum exactly: add/remove one interaction at a time from random users
    diff = int(total - deg.sum())
    for _guard in range(100000):
        if diff == 0:
            break
        step = 1 if diff > 0 else -1
        n = min(abs(diff), num_users)
        idx = rng.choice(num_users, size=n, replace=False)
        ok = (deg[idx] + step >= 1) & (deg[idx] + step <= cap)
        deg[idx[ok]] += step
        diff = int(total - deg.sum())
    if diff != 0:
        raise RuntimeError("degree fix-up did not converge")
    return deg


def generate(num_users: int, num_items: int, train_size: int, max_his: int = 100, seed: int = 2022,
             test_frac: float = 0.2, n_clusters: int = 64, p_in: float = 0.8, zipf_a: float = 1.0,
             max_degree: int | None = None, pop_offset: float | None = None) -> ClickData:
    """Item popularity is Zipf with a flattened head, p(rank) ~ 1/(rank + pop_offset)^zipf_a.  The default
    offset, max(10, num_items/300), puts the most popular item in ~2-3 % of the users' lists at the
    BASELINE shapes (Amazon-Books shape: ~1.4k of 2.38M train interactions), the order of magnitude of
    the real LightGCN datasets; offset 10 (the value small shapes get) would put one item in half of
    all Amazon-shape users' lists, which no real catalogue has and which makes Hogwild collisions on
    that row two orders of magnitude more frequent than in the reference's workloads."""
    rng = np.random.default_rng(seed)
    if pop_offset is None:
        pop_offset = max(10.0, num_items / 300.0)
    cap = max_degree if max_degree is not None else max(2, num_items // 8)
    cap = min(cap, int(num_items * (1.0 - test_frac)) - 1)
    deg_train = _degrees(rng, num_users, train_size, cap)
    deg_test = np.maximum(1, np.rint(deg_train * (test_frac / (1.0 - test_frac))).astype(np.int64))
    need = deg_train + deg_test

    n_clusters = max(1, min(n_clusters, num_items // 4))
    item_cluster = rng.integers(0, n_clusters, size=num_items)
    user_cluster = rng.integers(0, n_clusters, size=num_users)
    pop = 1.0 / (rng.permutation(num_items) + float(pop_offset)) ** zipf_a
    glob_cdf = np.cumsum(pop)
    order = np.argsort(item_cluster, kind="stable")          # items grouped by cluster
    clus_cdf = np.cumsum(pop[order])
    starts = np.searchsorted(item_cluster[order], np.arange(n_clusters), side="left")
    ends = np.searchsorted(item_cluster[order], np.arange(n_clusters), side="right")
    lo = np.where(starts > 0, clus_cdf[np.maximum(starts, 1) - 1], 0.0)
    hi = clus_cdf[np.maximum(ends, 1) - 1]

    chosen = [None] * num_users
    pending = np.arange(num_users)
    have = [np.empty(0, dtype=np.int64)] * num_users
    for _round in range(8):
        if pending.size == 0:
            break
        want = need[pending] - np.array([have[u].size for u in pending])
        m = (want * 3) // 2 + 8
        owner = np.repeat(pending, m)
        r = rng.random(owner.size)
        inside = rng.random(owner.size) < p_in
        c = user_cluster[owner]
        tgt = lo[c] + r * (hi[c] - lo[c])
        pos_in = np.minimum(np.searchsorted(clus_cdf, tgt, side="right"), num_items - 1)
        empty = ends[c] <= starts[c]
        item_in = order[np.clip(pos_in, starts[c], np.maximum(ends[c] - 1, starts[c]))]
        item_out = np.minimum(np.searchsorted(glob_cdf, r * glob_cdf[-1], side="right"), num_items - 1)
        items = np.where(inside & ~empty, item_in, item_out).astype(np.int64)
        off = np.concatenate([[0], np.cumsum(m)])
        nxt = []
        for j, u in enumerate(pending):
            cand = np.concatenate([have[u], items[off[j]:off[j + 1]]])
            _, first = np.unique(cand, return_index=True)
            cand = cand[np.sort(first)]
            if cand.size >= need[u]:
                chosen[u] = cand[:need[u]]
            else:
                have[u] = cand
                nxt.append(u)
        pending = np.array(nxt, dtype=np.int64)
    for u in pending:  # extremely unlikely; fill with unused ids deterministically
        used = set(have[u].tolist())
        extra = [i for i in range(num_items) if i not in used][: need[u] - have[u].size]
        chosen[u] = np.concatenate([have[u], np.array(extra, dtype=np.int64)])

    train_indptr = np.concatenate([[0], np.cumsum(deg_train)]).astype(np.int64)
    test_indptr = np.concatenate([[0], np.cumsum(deg_test)]).astype(np.int64)
    train_items = np.concatenate([chosen[u][:deg_train[u]] for u in range(num_users)]).astype(np.int32)
    test_items = np.concatenate([chosen[u][deg_train[u]:] for u in range(num_users)]).astype(np.int32)
    clicks = np.empty((train_size, 2), dtype=np.uint64)
    clicks[:, 0] = np.repeat(np.arange(num_users, dtype=np.uint64), deg_train)
    clicks[:, 1] = train_items.astype(np.uint64)
    his_items, masks = build_history(train_indptr, train_items, max_his, seed)
    return ClickData(num_users, num_items, max_his, clicks, his_items, masks,
                     train_indptr, train_items, test_indptr, test_items)


def build_history(train_indptr, train_items, max_his: int, seed: int):
    """cf_cpu/cf/datasets.py:47-61 with a seeded ``random`` (the reference leaves it unseeded)."""
    rnd = _pyrandom.Random(seed)
    num_users = train_indptr.size - 1
    his = np.zeros((num_users, max_his), dtype=np.uint64)
    masks = np.zeros((num_users, 1), dtype=np.uint64)
    for u in range(num_users):
        items = train_items[train_indptr[u]:train_indptr[u + 1]]
        n = items.size
        if n >= max_his:
            his[u] = rnd.sample(items.tolist(), max_his)
            masks[u, 0] = max_his
        elif n > 0:
            his[u, :n] = items
            his[u, n:] = items[-1]
            masks[u, 0] = n
    return his, masks


def generate_named(name: str, max_his: int = 100, seed: int = 2022, users_mult: int = 1) -> ClickData:
    """`users_mult` > 1 is the weak-scaling form of a shape: users_mult x the users and interactions over the SAME item
    catalogue (one seed, so one item-cluster map and one popularity law), to be split by user range across that many
    ranks (heat_b200/shard.py)."""
    u, i, n = SHAPES[name]
    return generate(u * users_mult, i, n * users_mult, max_his=max_his, seed=seed)


def init_tables(num_users: int, num_items: int, emb_dim: int, seed: int = 2022):
    """N(0, 1e-2^2) fp32 tables and W0 (models.py:15-16, behavior_aggregators.py:15), numpy-seeded."""
    rng = np.random.default_rng(seed + 1)
    user_w = (rng.standard_normal((num_users, emb_dim)) * 1e-2).astype(np.float32)
    item_w = (rng.standard_normal((num_items, emb_dim)) * 1e-2).astype(np.float32)
    w0 = (rng.standard_normal((emb_dim, emb_dim)) * 1e-2).astype(np.float32)
    return user_w, item_w, w0


def write_lightgcn_txt(data: ClickData, train_path: str, test_path: str) -> None:
    """Write ``user item item ...`` lines (the format cf_cpu/cf/datasets.py:31-40 parses)."""
    for path, indptr, items in ((train_path, data.train_indptr, data.train_items),
                                (test_path, data.test_indptr, data.test_items)):
        with open(path, "w") as f:
            for u in range(data.num_users):
                row = items[indptr[u]:indptr[u + 1]]
                f.write(" ".join([str(u)] + [str(int(x)) for x in row]) + "\n")


def generate_powerlaw(num_users: int, num_items: int, train_size: int, max_his: int = 100, seed: int = 2022,
                      test_frac: float = 0.2, deg_a: float = 0.6, zipf_a: float = 1.0, pop_offset: float | None = None,
                      max_degree: int = 4096) -> ClickData:
    """BASELINE.json configs[4]-style shard: power-law user degrees (~rank^-deg_a) and Zipf item popularity, no
    latent structure, fully vectorised (tens of millions of interactions in well under a minute).  Items are drawn
    with replacement and de-duplicated per user, so the train size comes out a few percent under `train_size`;
    a user's items are in random order, the first 1-test_frac of them are train (at least one), the history is the
    first max_his train items (a uniform sample, as datasets.py:47-61 draws it) padded with the last one."""
    rng = np.random.default_rng(seed)
    total = int(train_size / (1.0 - test_frac))
    raw = (rng.permutation(num_users) + 1.0) ** (-deg_a)
    deg = np.clip(np.rint(raw * (total / raw.sum())).astype(np.int64), 2, max_degree)
    if pop_offset is None:
        pop_offset = max(10.0, num_items / 300.0)
    cdf = np.cumsum(1.0 / (rng.permutation(num_items) + float(pop_offset)) ** zipf_a)
    owner = np.repeat(np.arange(num_users, dtype=np.int64), deg)
    # inverse CDF through a 4M-entry quantile table (a sorted searchsorted once, then gathers): exact enough for a
    # synthetic popularity law and ~20x faster than one binary search per draw
    T = 1 << 22
    table = np.minimum(np.searchsorted(cdf, (np.arange(T) + 0.5) * (cdf[-1] / T), side="right"), num_items - 1).astype(np.int64)
    item = table[rng.integers(0, T, size=owner.size)]
    key = np.sort((owner << 32) | item)
    key = key[np.concatenate([[True], key[1:] != key[:-1]])]            # drops a user's repeated items
    owner, item = key >> 32, (key & 0xffffffff).astype(np.int32)
    order = np.argsort((owner << 32) | rng.integers(0, 1 << 31, size=owner.size), kind="stable")   # random order inside every user
    owner, item = owner[order], item[order]
    n_u = np.bincount(owner, minlength=num_users).astype(np.int64)
    indptr = np.concatenate([[0], np.cumsum(n_u)])
    pos = np.arange(owner.size, dtype=np.int64) - indptr[owner]
    n_train = np.maximum(1, np.floor(n_u * (1.0 - test_frac)).astype(np.int64))
    n_train = np.minimum(n_train, n_u)
    is_train = pos < n_train[owner]
    train_items, test_items = item[is_train], item[~is_train]
    train_indptr = np.concatenate([[0], np.cumsum(n_train)]).astype(np.int64)
    test_indptr = np.concatenate([[0], np.cumsum(n_u - n_train)]).astype(np.int64)
    clicks = np.empty((train_items.size, 2), dtype=np.uint64)
    clicks[:, 0] = owner[is_train].astype(np.uint64)
    clicks[:, 1] = train_items.astype(np.uint64)
    h = np.minimum(n_train, max_his)
    j = np.minimum(np.arange(max_his, dtype=np.int64)[None, :], np.maximum(n_train, 1)[:, None] - 1)
    his = train_items[np.minimum(train_indptr[:-1, None] + j, max(train_items.size - 1, 0))].astype(np.uint64)
    his[n_train == 0] = 0
    masks = h.astype(np.uint64).reshape(-1, 1)
    return ClickData(num_users, num_items, max_his, clicks, his, masks, train_indptr, train_items, test_indptr, test_items)


def generate_powerlaw_block(block: int, users_per_block: int, num_items: int, samples_per_block: int, max_his: int = 100,
                            seed: int = 2022, test_frac: float = 0.2, deg_a: float = 0.6, zipf_a: float = 1.0,
                            n_clusters: int = 256, p_in: float = 0.8, max_degree: int = 4096) -> ClickData:
    """User block `block` of ONE large power-law dataset WITH planted structure (BASELINE.json configs[4] on 8 GPUs: eight
    blocks of 1.25 M users over one 2 M-item catalogue).  What all blocks share comes from `seed` alone -- the items'
    clusters and popularity law -- so the union of the blocks is one dataset whatever process generates which block; a
    block's users (power-law degrees, a cluster each, 80 % of their items drawn from that cluster by popularity) come
    from (seed, block).  Fully vectorised.  Local user ids 0..users_per_block-1; items global."""
    g = np.random.default_rng(seed)
    n_clusters = max(1, min(n_clusters, num_items // 8))
    item_cluster = g.integers(0, n_clusters, size=num_items)
    pop = 1.0 / (g.permutation(num_items) + max(10.0, num_items / 300.0)) ** zipf_a
    glob_cdf = np.cumsum(pop)
    order = np.argsort(item_cluster, kind="stable")
    clus_cdf = np.cumsum(pop[order])
    starts = np.searchsorted(item_cluster[order], np.arange(n_clusters), side="left")
    ends = np.searchsorted(item_cluster[order], np.arange(n_clusters), side="right")
    lo = np.where(starts > 0, clus_cdf[np.maximum(starts, 1) - 1], 0.0)
    hi = clus_cdf[np.maximum(ends, 1) - 1]

    rng = np.random.default_rng([seed, 7919, block])
    total = int(samples_per_block / (1.0 - test_frac))
    raw = (rng.permutation(users_per_block) + 1.0) ** (-deg_a)
    deg = np.clip(np.rint(raw * (total / raw.sum())).astype(np.int64), 2, max_degree)
    user_cluster = rng.integers(0, n_clusters, size=users_per_block)
    owner = np.repeat(np.arange(users_per_block, dtype=np.int64), deg)
    r = rng.random(owner.size)
    inside = rng.random(owner.size) < p_in
    c = user_cluster[owner]
    T = 1 << 22                                              # global draws through a quantile table (see generate_powerlaw)
    table = np.minimum(np.searchsorted(glob_cdf, (np.arange(T) + 0.5) * (glob_cdf[-1] / T), side="right"), num_items - 1)
    item = table[np.minimum((r * T).astype(np.int64), T - 1)]
    idx = np.nonzero(inside & (ends[c] > starts[c]))[0]
    ci = c[idx]
    pos_in = np.searchsorted(clus_cdf, lo[ci] + r[idx] * (hi[ci] - lo[ci]), side="right")
    item[idx] = order[np.clip(pos_in, starts[ci], ends[ci] - 1)]
    key = np.sort((owner << 32) | item.astype(np.int64))
    key = key[np.concatenate([[True], key[1:] != key[:-1]])]            # drops a user's repeated items
    owner, item = key >> 32, (key & 0xffffffff).astype(np.int32)
    shuffle = np.argsort((owner << 32) | rng.integers(0, 1 << 31, size=owner.size), kind="stable")   # random order inside every user
    owner, item = owner[shuffle], item[shuffle]
    n_u = np.bincount(owner, minlength=users_per_block).astype(np.int64)
    indptr = np.concatenate([[0], np.cumsum(n_u)])
    pos = np.arange(owner.size, dtype=np.int64) - indptr[owner]
    n_train = np.minimum(np.maximum(1, np.floor(n_u * (1.0 - test_frac)).astype(np.int64)), n_u)
    is_train = pos < n_train[owner]
    train_items, test_items = item[is_train], item[~is_train]
    train_indptr = np.concatenate([[0], np.cumsum(n_train)]).astype(np.int64)
    test_indptr = np.concatenate([[0], np.cumsum(n_u - n_train)]).astype(np.int64)
    clicks = np.empty((train_items.size, 2), dtype=np.uint64)
    clicks[:, 0] = owner[is_train].astype(np.uint64)
    clicks[:, 1] = train_items.astype(np.uint64)
    j = np.minimum(np.arange(max_his, dtype=np.int64)[None, :], np.maximum(n_train, 1)[:, None] - 1)
    his = train_items[np.minimum(train_indptr[:-1, None] + j, max(train_items.size - 1, 0))].astype(np.uint64)
    his[n_train == 0] = 0
    masks = np.minimum(n_train, max_his).astype(np.uint64).reshape(-1, 1)
    # the evaluation masks need every user's train items sorted by id; the click order stays random
    srt = np.argsort((owner[is_train] << 32) | train_items.astype(np.int64), kind="stable")
    return ClickData(users_per_block, num_items, max_his, clicks, his, masks, train_indptr, train_items[srt], test_indptr, test_items)


def concat_blocks(blocks) -> ClickData:
    """The union of user blocks as one ClickData (users renumbered block after block): what one process would train."""
    U = sum(b.num_users for b in blocks)
    clicks, off, so, to = [], 0, 0, 0
    tr_p, te_p = [np.zeros(1, np.int64)], [np.zeros(1, np.int64)]
    for b in blocks:
        c = b.clicks.copy(); c[:, 0] += np.uint64(off); clicks.append(c)
        tr_p.append(b.train_indptr[1:] + so); te_p.append(b.test_indptr[1:] + to)
        off += b.num_users; so += int(b.train_indptr[-1]); to += int(b.test_indptr[-1])
    return ClickData(U, blocks[0].num_items, blocks[0].max_his, np.concatenate(clicks), np.concatenate([b.his_items for b in blocks]),
                     np.concatenate([b.masks for b in blocks]), np.concatenate(tr_p), np.concatenate([b.train_items for b in blocks]),
                     np.concatenate(te_p), np.concatenate([b.test_items for b in blocks]))
