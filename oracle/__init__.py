"""CPU oracle for the HEAT MF_CCL hot path -- TEST INFRASTRUCTURE, not product code.

Two interchangeable back ends with one Python face:

* ``COracle``   -- ``oracle/heat_oracle.c``, a plain-C restatement (travels with the repo);
* ``RefOracle`` -- ``oracle/_ref/libheat_ref.so``, the UNMODIFIED reference sources compiled
  from /root/reference by ``oracle/Makefile`` (git-ignored build product, travels to the GPU
  box with the snapshot).  Use ``RefOracle.available()`` before constructing.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  ``heat_b200`` never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_u64p = C.POINTER(C.c_uint64)
_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)


def build(with_ref: bool = True) -> None:
    """Compile the C restatement and, if /root/reference is present, oracle/_ref."""
    subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"])
    if with_ref and os.path.isdir("/root/reference/cf_cpu/src"):
        subprocess.check_call(["make", "-s", "-C", _HERE, "ref"])


def _p(a, t):
    return a.ctypes.data_as(t)


def _check(a, dtype, name):
    if not (isinstance(a, np.ndarray) and a.dtype == dtype and a.flags.c_contiguous):
        raise TypeError(f"{name} must be a C-contiguous {dtype} array")


class _Base:
    """Shared face: borrows the caller's arrays and trains them IN PLACE."""

    def __init__(self, emb_dim, num_negs, clicks, his_items, masks, user_w, item_w, w0,
                 clip_val=1.0, l_r=1e-2, milestones=(10,), l2=1e-7):
        _check(clicks, np.uint64, "clicks"); _check(his_items, np.uint64, "his_items")
        _check(masks, np.uint64, "masks"); _check(user_w, np.float32, "user_w")
        _check(item_w, np.float32, "item_w"); _check(w0, np.float32, "w0")
        self.D, self.K = int(emb_dim), int(num_negs)
        self.U, self.I = int(user_w.shape[0]), int(item_w.shape[0])
        self.N, self.H = int(clicks.shape[0]), int(his_items.shape[1])
        self.clicks, self.his_items, self.masks = clicks, his_items, masks
        self.user_w, self.item_w, self.w0 = user_w, item_w, w0
        self.milestones = np.asarray(milestones, dtype=np.uint64)
        self.clip_val, self.l_r, self.l2 = float(clip_val), float(l_r), float(l2)
        self._h = None

    def grads(self):
        gu = np.empty((self.U, self.D), np.float32)
        gi = np.empty((self.I, self.D), np.float32)
        self._get_grads(self._h, _p(gu, _f32p), _p(gi, _f32p))
        return gu, gi

    def step(self, user, pos, neg_ids, reset_aggregator=False):
        negs = np.ascontiguousarray(neg_ids, dtype=np.uint64)
        assert negs.size == self.K
        return float(self._step(self._h, int(user), int(pos), _p(negs, _u64p), int(reset_aggregator)))

    def zero_grad(self):
        self._zero_grad(self._h)

    def set_lr(self, lr):
        self._set_lr(self._h, float(lr))

    def get_lr(self):
        return float(self._get_lr(self._h))

    def close(self):
        if self._h:
            self._destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class COracle(_Base):
    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            path = os.path.join(_HERE, "libheat_oracle.so")
            if not os.path.exists(path):
                build(with_ref=False)
            L = C.CDLL(path)
            L.heat_oracle_create.restype = C.c_void_p
            L.heat_oracle_create.argtypes = [C.c_uint64] * 6 + [C.c_float, C.c_float, _u64p, C.c_uint64,
                                                                _u64p, _u64p, _u64p, _f32p, _f32p, _f32p]
            L.heat_oracle_destroy.argtypes = [C.c_void_p]
            L.heat_oracle_get_grads.argtypes = [C.c_void_p, _f32p, _f32p]
            L.heat_oracle_zero_grad.argtypes = [C.c_void_p]
            L.heat_oracle_set_lr.argtypes = [C.c_void_p, C.c_float]
            L.heat_oracle_get_lr.argtypes = [C.c_void_p]; L.heat_oracle_get_lr.restype = C.c_float
            L.heat_oracle_step.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, _u64p, C.c_int]
            L.heat_oracle_step.restype = C.c_float
            L.heat_oracle_train_epoch.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int]
            L.heat_oracle_train_epoch.restype = C.c_double
            L.heat_oracle_neg_table_tile.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, _u64p, C.c_uint64, C.c_uint64, C.c_uint64,
                                                     C.c_uint64, C.c_uint64, _u64p]
            L.heat_oracle_neg_table.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, _u64p, C.c_uint64,
                                                C.c_uint64, C.c_uint64, _u64p]
            L.heat_oracle_philox.argtypes = [C.POINTER(C.c_uint32)] * 3
            L.heat_oracle_scores.argtypes = [C.c_void_p, _f32p]
            L.heat_oracle_topk_row.argtypes = [_f32p, C.c_uint64, _i32p, C.c_uint64, C.c_uint64, _i32p, _f32p]
            L.heat_oracle_metrics_row.argtypes = [_i32p, C.c_uint64, _i32p, C.c_uint64, C.POINTER(C.c_double)]
            cls._lib = L
        return cls._lib

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        L = self.lib()
        self._h = L.heat_oracle_create(self.D, self.K, self.U, self.I, self.N, self.H, self.clip_val, self.l_r,
                                       _p(self.milestones, _u64p), self.milestones.size,
                                       _p(self.clicks, _u64p), _p(self.his_items, _u64p), _p(self.masks, _u64p),
                                       _p(self.user_w, _f32p), _p(self.item_w, _f32p), _p(self.w0, _f32p))
        self._get_grads, self._step, self._zero_grad = L.heat_oracle_get_grads, L.heat_oracle_step, L.heat_oracle_zero_grad
        self._set_lr, self._get_lr, self._destroy = L.heat_oracle_set_lr, L.heat_oracle_get_lr, L.heat_oracle_destroy

    def train_epoch(self, seed=2022, sample_offset=0, nthreads=1):
        return float(self.lib().heat_oracle_train_epoch(self._h, int(seed), int(sample_offset), int(nthreads)))

    def scores(self):
        out = np.empty((self.U, self.I), np.float32)
        self.lib().heat_oracle_scores(self._h, _p(out, _f32p))
        return out


def philox4x32_10(ctr, key):
    c = (C.c_uint32 * 4)(*[int(x) & 0xFFFFFFFF for x in ctr])
    k = (C.c_uint32 * 2)(*[int(x) & 0xFFFFFFFF for x in key])
    o = (C.c_uint32 * 4)()
    COracle.lib().heat_oracle_philox(c, k, o)
    return [int(x) for x in o]


def neg_table(seed, epoch, sample_offset, clicks, num_negs, num_items, tile_size=0, refresh_interval=1):
    """[N, K] uint64 negative ids of one epoch under the Philox contract (oracle/philox_ref.h); tile_size > 0 =
    the hot-tile mapping."""
    _check(clicks, np.uint64, "clicks")
    out = np.empty((clicks.shape[0], num_negs), np.uint64)
    if tile_size:
        COracle.lib().heat_oracle_neg_table_tile(int(seed), int(epoch), int(sample_offset), _p(clicks, _u64p), clicks.shape[0],
                                                 int(num_negs), int(num_items), int(tile_size), int(refresh_interval), _p(out, _u64p))
        return out
    COracle.lib().heat_oracle_neg_table(int(seed), int(epoch), int(sample_offset), _p(clicks, _u64p),
                                        clicks.shape[0], int(num_negs), int(num_items), _p(out, _u64p))
    return out


def topk_and_metrics(scores, train_indptr, train_items, test_indptr, test_items, ks=(20, 50)):
    """cf/metrics.py evaluate_metrics restated: per-user masked top-max(ks), then mean
    Recall/NDCG/HitRate at each k over users with at least one test item.
    Returns (topk_ids [U,max_k] int32, topk_scores, {"Recall(k=20)": ..., ...})."""
    L = COracle.lib()
    U, I = scores.shape
    max_k = max(ks)
    ids = np.empty((U, max_k), np.int32)
    sc = np.empty((U, max_k), np.float32)
    train_items = np.ascontiguousarray(train_items, np.int32)
    test_items = np.ascontiguousarray(test_items, np.int32)
    sums = {k: np.zeros(3) for k in ks}
    n_eval = 0
    out3 = (C.c_double * 3)()
    for u in range(U):
        row = np.ascontiguousarray(scores[u]).copy()
        a, b = int(train_indptr[u]), int(train_indptr[u + 1])
        tr = train_items[a:b]
        L.heat_oracle_topk_row(_p(row, _f32p), I, _p(tr, _i32p) if b > a else None, b - a, max_k,
                               _p(ids[u], _i32p), _p(sc[u], _f32p))
        a, b = int(test_indptr[u]), int(test_indptr[u + 1])
        if b > a:
            te = test_items[a:b]
            n_eval += 1
            for k in ks:
                L.heat_oracle_metrics_row(_p(ids[u], _i32p), k, _p(te, _i32p), b - a, out3)
                sums[k] += np.array(out3[:])
    res = {}
    for k in ks:
        m = sums[k] / max(n_eval, 1)
        res[f"Recall(k={k})"], res[f"NDCG(k={k})"], res[f"HitRate(k={k})"] = map(float, m)
    return ids, sc, res


class RefOracle(_Base):
    _libs = {}

    @staticmethod
    def path(opt="O3"):
        return os.path.join(_HERE, "_ref", "libheat_ref.so" if opt == "O3" else "libheat_ref_O0.so")

    @classmethod
    def available(cls, opt="O3"):
        return os.path.exists(cls.path(opt))

    @classmethod
    def lib(cls, opt="O3"):
        if opt not in cls._libs:
            L = C.CDLL(cls.path(opt))
            L.heat_ref_create.restype = C.c_void_p
            L.heat_ref_create.argtypes = [C.c_uint64] * 6 + [C.c_float] * 3 + [_u64p, C.c_uint64,
                                                                               _u64p, _u64p, _u64p, _f32p, _f32p, _f32p]
            L.heat_ref_destroy.argtypes = [C.c_void_p]
            L.heat_ref_get_grads.argtypes = [C.c_void_p, _f32p, _f32p]
            L.heat_ref_zero_grad.argtypes = [C.c_void_p]
            L.heat_ref_set_lr.argtypes = [C.c_void_p, C.c_float]
            L.heat_ref_get_lr.argtypes = [C.c_void_p]; L.heat_ref_get_lr.restype = C.c_float
            L.heat_ref_step.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, _u64p, C.c_int]
            L.heat_ref_step.restype = C.c_float
            L.heat_ref_train_epoch.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_uint64, C.c_int]
            L.heat_ref_train_epoch.restype = C.c_double
            L.heat_ref_evaluate0.argtypes = [C.c_void_p, _f32p]
            L.heat_ref_max_threads.restype = C.c_int
            cls._libs[opt] = L
        return cls._libs[opt]

    def __init__(self, *a, opt="O3", **kw):
        super().__init__(*a, **kw)
        L = self._L = self.lib(opt)
        self._h = L.heat_ref_create(self.D, self.K, self.U, self.I, self.N, self.H, self.l2, self.clip_val, self.l_r,
                                    _p(self.milestones, _u64p), self.milestones.size,
                                    _p(self.clicks, _u64p), _p(self.his_items, _u64p), _p(self.masks, _u64p),
                                    _p(self.user_w, _f32p), _p(self.item_w, _f32p), _p(self.w0, _f32p))
        self._get_grads, self._step, self._zero_grad = L.heat_ref_get_grads, L.heat_ref_step, L.heat_ref_zero_grad
        self._set_lr, self._get_lr, self._destroy = L.heat_ref_set_lr, L.heat_ref_get_lr, L.heat_ref_destroy

    def train_epoch(self, seed=2022, sample_offset=0, nthreads=1, sampler=2):
        """sampler 0 = reference mt19937 sampler; 1 = Philox through the reference's persistent
        ignore_pos vector (deterministic only at nthreads=1); 2 = Philox stateless look-back."""
        return float(self._L.heat_ref_train_epoch(self._h, int(sampler), int(seed), int(sample_offset), int(nthreads)))

    def scores(self):
        out = np.empty((self.U, self.I), np.float32)
        self._L.heat_ref_evaluate0(self._h, _p(out, _f32p))
        return out

    def max_threads(self):
        return int(self._L.heat_ref_max_threads())
