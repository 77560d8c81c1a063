"""LightGCN-format click data (reference: cf_cpu/cf/datasets.py:14-216).

File format (datasets.py:31-40): one line per user, `user item item ...`.  Arrays handed to cf_c
(the reference's layout): click_dataset uint64 [N, 2] in file order, his_items uint64
[num_users, max_his], masks uint64 [num_users, 1].  History rule (datasets.py:47-61): a user with
>= max_his items gets a random sample of max_his of them, otherwise all items padded with the last
one; mask = min(len, max_his).  The reference samples with an unseeded `random`; here the RNG is
seeded from the config (`seed`, default 2022) so runs are reproducible.

Differences from the reference, both deliberate: the file is parsed natively (`cf_c.modules.datasets.parse_clicks`
-> heat_clickfile_*, one thread per core) into a CSR (indptr, items) instead of python lists of pairs built line
by line, the CSR-by-user and history arrays are built with vectorised numpy (python only samples the users that
have more than max_his items), and a rank takes its user range by slicing (`SubClickDataset`) without rank 0
pickling shards to the others (main.py:55-69).
"""
import random

import numpy as np

try:
    from .cpp_base import CPPBase
    from . import cf_c
except ImportError:
    from cpp_base import CPPBase
    import cf_c


class Dataset(CPPBase):
    def __init__(self):
        super().__init__()


def _history(indptr, items, max_his, seed):
    """datasets.py:47-61.  Users below max_his items (the bulk) are filled by one gather; only the users with
    >= max_his items go through python's `random.sample`, in user order (the seeded stream is unchanged)."""
    rnd = random.Random(seed)
    num_users = indptr.size - 1
    n = np.diff(indptr).astype(np.int64)
    masks = np.minimum(n, max_his).astype(np.uint64).reshape(-1, 1)
    if items.size:
        j = np.minimum(np.arange(max_his, dtype=np.int64)[None, :], np.maximum(n, 1)[:, None] - 1)
        his = items[np.minimum(indptr[:-1, None] + j, items.size - 1)].astype(np.uint64)
        his[n == 0] = 0
    else:
        his = np.zeros((num_users, max_his), dtype=np.uint64)
    for u in np.nonzero(n >= max_his)[0]:
        his[u] = rnd.sample(items[indptr[u]:indptr[u + 1]].tolist(), max_his)
    for u in np.nonzero(n == 0)[0][:20]:
        print(f"Warning {u} has 0 items !!! ")
    return his, masks


def parse_python(file_path, separator=' '):
    """The reference's line-by-line reader (datasets.py:31-40), kept as the checker of the native one."""
    users, rows = [], []
    with open(file_path, mode='r') as f:
        for line in f:
            line = line.strip()
            if not line:
                continue
            if separator != ' ':
                line = line.replace(separator, ' ')
            vals = np.array(line.split(), dtype=np.int64)
            users.append(int(vals[0]))
            rows.append(vals[1:])
    counts = np.array([r.size for r in rows], dtype=np.int64)
    indptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    items = np.concatenate(rows).astype(np.int64) if rows else np.zeros(0, np.int64)
    return np.asarray(users, dtype=np.int64), indptr, items


class ClickDataset(Dataset):
    def __init__(self, file_path, separator=' ', config=None):
        super().__init__()
        self.c_class = cf_c.modules.datasets.ClickDataset
        self.file_path = file_path
        self.max_his = config.max_his
        self.rank = 0
        users, file_indptr, file_items = cf_c.modules.datasets.parse_clicks(file_path, separator)
        self.num_users = int(users.size)
        if self.num_users and (users.min() != 0 or users.max() != self.num_users - 1 or np.unique(users).size != self.num_users):
            print('Warning user_id is not continuous! ')
        # rows are stored at their user id, samples stay in FILE order (engine.cpp:330-331 walks them that way)
        self._file_users, self._file_indptr, self._file_items = users, file_indptr, file_items
        self.num_items = int(np.unique(self._file_items).size)
        self._finish(config, is_train='train' in file_path)

    def _finish(self, config, is_train):
        n_rows = max(self.num_users, int(self._file_users.max()) + 1 if self._file_users.size else 0)
        per_user = np.bincount(self._file_users, weights=np.diff(self._file_indptr), minlength=n_rows).astype(np.int64)
        self.indptr = np.concatenate([[0], np.cumsum(per_user)]).astype(np.int64)
        # CSR by user id: a stable sort of the file's (user, item) pairs by user (identity when the file is sorted)
        pair_user = np.repeat(self._file_users, np.diff(self._file_indptr))
        if self._file_users.size and np.all(self._file_users[1:] > self._file_users[:-1]):
            self.items = self._file_items.astype(np.int32)
        else:
            self.items = self._file_items[np.argsort(pair_user, kind="stable")].astype(np.int32)
        self.click_dataset = np.empty((self._file_items.size, 2), dtype=np.uint64)
        self.click_dataset[:, 0] = pair_user.astype(np.uint64)
        self.click_dataset[:, 1] = self._file_items.astype(np.uint64)
        self.his_items, self.masks = _history(self.indptr, self.items, self.max_his, getattr(config, "seed", 2022))
        self.gen_dataset_info()
        if is_train:
            self.update_config(config)

    # -- reference-compatible views -------------------------------------------------------------
    @property
    def user_items_dic(self):
        return {u: self.items[self.indptr[u]:self.indptr[u + 1]].tolist() for u in range(self.indptr.size - 1)
                if self.indptr[u + 1] > self.indptr[u]}

    @property
    def user_item_ids(self):
        return self.click_dataset.tolist()

    def get_user_items(self):
        return self.user_items_dic

    def csr(self, num_users=None):
        """(indptr int64 [num_users+1], items int32) -- the form the fused evaluation takes."""
        indptr = self.indptr
        if num_users is not None and indptr.size - 1 != num_users:
            if indptr.size - 1 > num_users:
                raise ValueError("dataset has more users than the model")
            indptr = np.concatenate([indptr, np.full(num_users + 1 - indptr.size, indptr[-1])])
        return np.ascontiguousarray(indptr, np.int64), np.ascontiguousarray(self.items, np.int32)

    def gen_dataset_info(self):
        print(f'gen dataset info of {self.file_path} ')
        print(f'number of users: {self.num_users}; number of items: {self.num_items}')
        if self.items.size and int(self.items.max()) + 1 != self.num_items:
            print('Warning item_id is not continuous! ')
        print(f'total samples: {self.click_dataset.shape[0]} ')

    def update_config(self, config=None):
        """datasets.py:73-79 / 200-212: back-fill the dataset-derived config fields, build the cf_c object."""
        print('update config, init c_instance !!! ')
        config.num_users = self.num_users
        config.num_items = max(self.num_items, int(self.items.max()) + 1 if self.items.size else 0)
        config.train_size = int(self.click_dataset.shape[0])
        self.init_c_instance(click_dataset=self.click_dataset, historical_items=self.his_items, masks=self.masks)
        self.c_instance.max_his = self.max_his


class SubClickDataset(ClickDataset):
    """Users [start, end) of a parent dataset with user ids re-based to 0 (datasets.py:110-158).
    History rows are sliced from the parent (the reference re-samples them, unseeded)."""

    def __init__(self, parent_dataset, start, end, rank=None):
        Dataset.__init__(self)
        self.c_class = cf_c.modules.datasets.ClickDataset
        p = parent_dataset
        self.file_path, self.max_his, self.rank = p.file_path, p.max_his, rank
        end = min(end, p.indptr.size - 1)
        s0, s1 = int(p.indptr[start]), int(p.indptr[end])
        self.indptr = (p.indptr[start:end + 1] - s0).astype(np.int64)
        self.items = p.items[s0:s1].copy()
        sel = (p.click_dataset[:, 0] >= start) & (p.click_dataset[:, 0] < end)
        self.click_dataset = np.ascontiguousarray(p.click_dataset[sel])
        self.click_dataset[:, 0] -= np.uint64(start)
        self.sample_offset = int(np.argmax(sel)) if sel.any() else 0      # global index of the first local sample
        self.his_items = np.ascontiguousarray(p.his_items[start:end])
        self.masks = np.ascontiguousarray(p.masks[start:end])
        self.num_users = end - start
        self.num_items = p.num_items
        self._parent_num_items = max(p.num_items, int(p.items.max()) + 1 if p.items.size else 0)
        print(f'number of users: {self.num_users}; number of items: {self.num_items}; total samples: {self.click_dataset.shape[0]} ')

    def update_config(self, config=None):
        super().update_config(config)
        config.num_items = self._parent_num_items          # item ids stay global
        config.sample_offset = self.sample_offset
