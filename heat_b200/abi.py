"""ctypes face of the C ABI (include/heat_b200.h) -- used by bench.py and the GPU tests.

The reference-facing plugin is the pybind11 module ``cf_c`` (heat_b200/cf/); this module
is the thin, dependency-free way to reach the same entry points from Python.  There is no
fallback: a missing library or a failing call raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libheat_b200.so")

_u64p = C.POINTER(C.c_uint64)
_i64p = C.POINTER(C.c_int64)
_i32p = C.POINTER(C.c_int32)
_f32p = C.POINTER(C.c_float)
_f64p = C.POINTER(C.c_double)
_u8p = C.POINTER(C.c_uint8)

# every symbol include/heat_b200.h declares (tests check the library exports them all)
SYMBOLS = [
    "heat_abi_version", "heat_last_error", "heat_create", "heat_destroy", "heat_set_dataset",
    "heat_bind_tables", "heat_upload_tables", "heat_sync_to_host", "heat_train_epoch", "heat_train_range",
    "heat_zero_grad", "heat_get_grads", "heat_set_lr", "heat_get_lr", "heat_get_epoch", "heat_set_epoch",
    "heat_last_kernel_ms", "heat_last_epoch_ms", "heat_launch_count", "heat_sample_negatives", "heat_history_indices",
    "heat_philox_draw", "heat_eval_dense", "heat_eval_topk", "heat_eval_metrics", "heat_last_eval_ms",
    "heat_comm_unique_id", "heat_comm_init", "heat_comm_sync_items", "heat_set_sync_interval", "heat_comm_allreduce_f64",
    "heat_probe_l2_bandwidth", "heat_host_alloc", "heat_host_free", "heat_save_state", "heat_load_state", "heat_clickfile_open", "heat_clickfile_read", "heat_clickfile_close", "heat_tile_draw", "heat_last_engine", "heat_comm_mode",
    "heat_train_range_noalloc", "heat_last_exchange", "heat_local_group_init", "heat_local_group_train_epoch",
    "heat_local_group_share", "heat_probe_tf32_peak",
]


class HeatConfig(C.Structure):
    _fields_ = [
        ("emb_dim", C.c_uint64), ("num_negs", C.c_uint64), ("num_users", C.c_uint64), ("num_items", C.c_uint64),
        ("train_size", C.c_uint64), ("neg_sampler", C.c_uint64), ("tile_size", C.c_uint64),
        ("refresh_interval", C.c_uint64), ("num_subepochs", C.c_uint64),
        ("l2", C.c_float), ("clip_val", C.c_float), ("l_r", C.c_float),
        ("milestones", _u64p), ("n_milestones", C.c_uint64),
        ("seed", C.c_uint64), ("sample_offset", C.c_uint64),
        ("chunk", C.c_uint32), ("deterministic", C.c_uint32), ("w0_tail_mode", C.c_uint32),
        ("warps_per_block", C.c_uint32), ("blocks_per_sm", C.c_uint32), ("update_mode", C.c_uint32),
        ("sync_average", C.c_uint32), ("history_mode", C.c_uint32), ("engine", C.c_uint32), ("max_blocks", C.c_uint32),
        ("exchange_mode", C.c_uint32), ("tile_sampler", C.c_uint32), ("sync_grads", C.c_uint32), ("eval_precision", C.c_uint32),
    ]


NOALLOC_MAX_SAMPLES = 1 << 20      # HEAT_NOALLOC_MAX_SAMPLES


class HeatError(RuntimeError):
    pass


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise HeatError(f"{LIB_PATH} is missing: run `python -m heat_b200.build` (no CPU fallback exists)")
        L = C.CDLL(LIB_PATH)
        L.heat_last_error.restype = C.c_char_p
        L.heat_create.argtypes = [C.POINTER(C.c_void_p), C.POINTER(HeatConfig), C.c_int]
        L.heat_destroy.argtypes = [C.c_void_p]
        L.heat_set_dataset.argtypes = [C.c_void_p, _u64p, C.c_uint64, _u64p, C.c_uint64, C.c_uint64, _u64p]
        L.heat_bind_tables.argtypes = [C.c_void_p, _f32p, _f32p, _f32p]
        L.heat_upload_tables.argtypes = [C.c_void_p]
        L.heat_sync_to_host.argtypes = [C.c_void_p]
        L.heat_train_epoch.argtypes = [C.c_void_p, _f64p]
        L.heat_train_range.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, _f64p, _f32p]
        L.heat_train_range_noalloc.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, _f64p, _f32p]
        L.heat_zero_grad.argtypes = [C.c_void_p]
        L.heat_get_grads.argtypes = [C.c_void_p, _f32p, _f32p]
        L.heat_clickfile_open.argtypes = [C.c_char_p, C.c_char, C.POINTER(C.c_void_p), _u64p, _u64p]
        L.heat_clickfile_read.argtypes = [C.c_void_p, _i64p, _i64p, _i64p]
        L.heat_clickfile_close.argtypes = [C.c_void_p]
        L.heat_save_state.argtypes = [C.c_void_p, C.c_char_p]
        L.heat_load_state.argtypes = [C.c_void_p, C.c_char_p]
        L.heat_set_lr.argtypes = [C.c_void_p, C.c_float]
        L.heat_get_lr.argtypes = [C.c_void_p, _f32p]
        L.heat_get_epoch.argtypes = [C.c_void_p, _u64p]
        L.heat_set_epoch.argtypes = [C.c_void_p, C.c_uint64]
        L.heat_last_kernel_ms.argtypes = [C.c_void_p, _f32p]
        L.heat_last_eval_ms.argtypes = [C.c_void_p, _f32p]
        L.heat_last_epoch_ms.argtypes = [C.c_void_p, _f32p]
        L.heat_launch_count.argtypes = [C.c_void_p, _u64p]
        L.heat_last_engine.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        L.heat_sample_negatives.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, _u64p]
        L.heat_history_indices.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, _u64p, _u64p]
        L.heat_philox_draw.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint64]
        L.heat_philox_draw.restype = C.c_uint64
        L.heat_tile_draw.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint64]
        L.heat_tile_draw.restype = C.c_uint64
        L.heat_eval_dense.argtypes = [C.c_void_p, _f32p]
        L.heat_eval_topk.argtypes = [C.c_void_p, C.c_uint32, _i64p, _i32p, _i32p, _f32p]
        L.heat_eval_metrics.argtypes = [C.c_void_p, _i64p, _i32p, _i64p, _i32p, _f64p, _f64p]
        L.heat_comm_unique_id.argtypes = [_u8p]
        L.heat_comm_init.argtypes = [C.c_void_p, C.c_int, C.c_int, _u8p]
        L.heat_comm_sync_items.argtypes = [C.c_void_p]
        L.heat_comm_mode.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        L.heat_set_sync_interval.argtypes = [C.c_void_p, C.c_uint64]
        L.heat_last_exchange.argtypes = [C.c_void_p, _u64p, _f32p]
        L.heat_local_group_init.argtypes = [C.POINTER(C.c_void_p), C.c_int]
        L.heat_local_group_train_epoch.argtypes = [C.POINTER(C.c_void_p), C.c_int, _f64p]
        L.heat_local_group_share.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_uint32]
        L.heat_comm_allreduce_f64.argtypes = [C.c_void_p, _f64p, C.c_int]
        L.heat_probe_l2_bandwidth.argtypes = [C.c_int, C.c_uint64, C.c_int, _f64p]
        L.heat_probe_tf32_peak.argtypes = [C.c_int, C.c_int, _f64p]
        L.heat_host_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_uint64]
        L.heat_host_free.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def _ck(rc):
    if rc != 0:
        raise HeatError(lib().heat_last_error().decode("utf-8", "replace"))


def _ptr(a, t):
    return a.ctypes.data_as(t) if a is not None else None


def _need(a, dtype, name):
    if not (isinstance(a, np.ndarray) and a.dtype == dtype and a.flags.c_contiguous):
        raise TypeError(f"{name} must be a C-contiguous {np.dtype(dtype).name} ndarray")
    return a


_pinned_keepalive = []


def pinned_empty(shape, dtype):
    """numpy array over page-locked host memory (heat_host_alloc); lives until process exit."""
    dtype = np.dtype(dtype)
    count = int(np.prod(shape))
    n = max(count * dtype.itemsize, 1)
    p = C.c_void_p()
    _ck(lib().heat_host_alloc(C.byref(p), n))
    buf = (C.c_uint8 * n).from_address(p.value)
    _pinned_keepalive.append((p, buf))
    return np.frombuffer(buf, dtype=dtype, count=count).reshape(shape)


def probe_l2_bandwidth(device=0, nbytes=74 << 20, iters=200):
    g = C.c_double()
    _ck(lib().heat_probe_l2_bandwidth(device, nbytes, iters, C.byref(g)))
    return g.value


def probe_tf32_peak(device=0, iters=4000):
    """Measured dense TF32 tcgen05.mma peak of the device, TFLOP/s (heat_probe_tf32_peak)."""
    g = C.c_double()
    _ck(lib().heat_probe_tf32_peak(device, iters, C.byref(g)))
    return g.value


def parse_clickfile(path, separator=" "):
    """LightGCN-format file -> (row_users int64 [R], row_indptr int64 [R+1], items int64 [nnz]) in file order,
    parsed natively with one thread per core (heat_clickfile_*; host code, needs no GPU)."""
    h, rows, nnz = C.c_void_p(), C.c_uint64(), C.c_uint64()
    sep = separator.encode()[:1] or b" "
    _ck(lib().heat_clickfile_open(os.fsencode(path), sep, C.byref(h), C.byref(rows), C.byref(nnz)))
    try:
        users = np.empty(rows.value, np.int64); indptr = np.empty(rows.value + 1, np.int64); items = np.empty(nnz.value, np.int64)
        _ck(lib().heat_clickfile_read(h, _ptr(users, _i64p), _ptr(indptr, _i64p), _ptr(items, _i64p)))
    finally:
        lib().heat_clickfile_close(h)
    return users, indptr, items


def tile_draw(seed, epoch, sample, k, num_items, tile_size, refresh_interval):
    """Host twin of the device sampler with the hot-tile mapping (tile_size 0 = uniform)."""
    return int(lib().heat_tile_draw(seed, epoch, sample, k, num_items, tile_size, refresh_interval))


def philox_draw(seed, epoch, sample, k, num_items):
    return int(lib().heat_philox_draw(seed, epoch, sample, k, num_items))


class Context:
    """One device-resident training/eval context (one per GPU)."""

    def __init__(self, emb_dim, num_negs, num_users, num_items, clip_val=1.0, l_r=1e-2, milestones=(10,),
                 l2=1e-7, seed=2022, sample_offset=0, chunk=0, deterministic=False, w0_tail_mode=0,
                 warps_per_block=0, blocks_per_sm=0, max_blocks=0, update_mode=0, sync_average=0, history_mode=0, engine=0, neg_sampler=0, tile_size=512, refresh_interval=8192,
                 tile_sampler=0, exchange_mode=0, sync_grads=0, eval_precision=0, device=0):
        self._ms = np.asarray(list(milestones), dtype=np.uint64)
        cfg = HeatConfig(emb_dim=emb_dim, num_negs=num_negs, num_users=num_users, num_items=num_items,
                         train_size=0, neg_sampler=neg_sampler, tile_size=tile_size,
                         refresh_interval=refresh_interval, num_subepochs=2, l2=l2, clip_val=clip_val, l_r=l_r,
                         milestones=_ptr(self._ms, _u64p), n_milestones=self._ms.size, seed=seed,
                         sample_offset=sample_offset, chunk=chunk, deterministic=int(bool(deterministic)),
                         w0_tail_mode=w0_tail_mode, warps_per_block=warps_per_block, blocks_per_sm=blocks_per_sm,
                         max_blocks=max_blocks, update_mode=update_mode,
                         sync_average=sync_average, history_mode=history_mode, engine=engine, tile_sampler=tile_sampler,
                         exchange_mode=exchange_mode, sync_grads=sync_grads, eval_precision=eval_precision)
        self.D, self.K, self.U, self.I = emb_dim, num_negs, num_users, num_items
        self.N = 0
        self.H = 0
        self._h = C.c_void_p()
        self._keep = []
        _ck(lib().heat_create(C.byref(self._h), C.byref(cfg), device))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().heat_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- inputs
    def set_dataset(self, clicks, his_items, masks):
        _need(clicks, np.uint64, "clicks"); _need(his_items, np.uint64, "his_items"); _need(masks, np.uint64, "masks")
        self.N, self.H = int(clicks.shape[0]), int(his_items.shape[1])
        _ck(lib().heat_set_dataset(self._h, _ptr(clicks, _u64p), self.N, _ptr(his_items, _u64p),
                                   his_items.shape[0], self.H, _ptr(masks, _u64p)))

    def bind_tables(self, user_w, item_w, w0):
        _need(user_w, np.float32, "user_w"); _need(item_w, np.float32, "item_w"); _need(w0, np.float32, "w0")
        assert user_w.shape == (self.U, self.D) and item_w.shape == (self.I, self.D) and w0.shape == (self.D, self.D)
        self._keep = [user_w, item_w, w0]
        _ck(lib().heat_bind_tables(self._h, _ptr(user_w, _f32p), _ptr(item_w, _f32p), _ptr(w0, _f32p)))

    def upload_tables(self):
        _ck(lib().heat_upload_tables(self._h))

    def sync_to_host(self):
        _ck(lib().heat_sync_to_host(self._h))

    # -- training
    def train_epoch(self):
        loss = C.c_double()
        _ck(lib().heat_train_epoch(self._h, C.byref(loss)))
        return loss.value

    def train_range(self, begin, end, per_sample=False):
        loss = C.c_double()
        out = np.empty(end - begin, np.float32) if per_sample else None
        _ck(lib().heat_train_range(self._h, begin, end, C.byref(loss), _ptr(out, _f32p)))
        return (loss.value, out) if per_sample else loss.value

    def train_range_noalloc(self, begin, end, per_sample=False):
        """heat_train_range without any device allocation (static scratch, <= NOALLOC_MAX_SAMPLES samples per call)."""
        loss = C.c_double()
        out = np.empty(end - begin, np.float32) if per_sample else None
        _ck(lib().heat_train_range_noalloc(self._h, begin, end, C.byref(loss), _ptr(out, _f32p)))
        return (loss.value, out) if per_sample else loss.value

    def zero_grad(self):
        _ck(lib().heat_zero_grad(self._h))

    def grads(self):
        gu = np.empty((self.U, self.D), np.float32); gi = np.empty((self.I, self.D), np.float32)
        _ck(lib().heat_get_grads(self._h, _ptr(gu, _f32p), _ptr(gi, _f32p)))
        return gu, gi

    def save_state(self, path):
        """Checkpoint {tables, grad rows, W0, epoch, LR} from the device to one file."""
        _ck(lib().heat_save_state(self._h, os.fsencode(path)))

    def load_state(self, path):
        _ck(lib().heat_load_state(self._h, os.fsencode(path)))

    def set_lr(self, lr):
        _ck(lib().heat_set_lr(self._h, lr))

    def get_lr(self):
        v = C.c_float(); _ck(lib().heat_get_lr(self._h, C.byref(v))); return v.value

    def get_epoch(self):
        v = C.c_uint64(); _ck(lib().heat_get_epoch(self._h, C.byref(v))); return v.value

    def set_epoch(self, e):
        _ck(lib().heat_set_epoch(self._h, e))

    def last_kernel_ms(self):
        v = C.c_float(); _ck(lib().heat_last_kernel_ms(self._h, C.byref(v))); return v.value

    def last_epoch_ms(self):
        v = C.c_float(); _ck(lib().heat_last_epoch_ms(self._h, C.byref(v))); return v.value

    def last_eval_ms(self):
        v = C.c_float(); _ck(lib().heat_last_eval_ms(self._h, C.byref(v))); return v.value

    def last_engine(self):
        v = C.c_int(); _ck(lib().heat_last_engine(self._h, C.byref(v))); return v.value

    def launch_count(self):
        v = C.c_uint64(); _ck(lib().heat_launch_count(self._h, C.byref(v))); return v.value

    # -- index contracts
    def sample_negatives(self, epoch, begin, end):
        out = np.empty((end - begin, self.K), np.uint64)
        _ck(lib().heat_sample_negatives(self._h, epoch, begin, end, _ptr(out, _u64p)))
        return out

    def history_indices(self, begin, end):
        ids = np.empty((end - begin, self.H), np.uint64); cnt = np.empty(end - begin, np.uint64)
        _ck(lib().heat_history_indices(self._h, begin, end, _ptr(ids, _u64p), _ptr(cnt, _u64p)))
        return ids, cnt

    # -- evaluation
    def eval_dense(self):
        out = np.empty((self.U, self.I), np.float32)
        _ck(lib().heat_eval_dense(self._h, _ptr(out, _f32p)))
        return out

    def eval_topk(self, k, train_indptr=None, train_items=None):
        ids = np.empty((self.U, k), np.int32); sc = np.empty((self.U, k), np.float32)
        if train_indptr is not None:
            train_indptr = np.ascontiguousarray(train_indptr, np.int64)
            train_items = np.ascontiguousarray(train_items, np.int32)
        _ck(lib().heat_eval_topk(self._h, k, _ptr(train_indptr, _i64p), _ptr(train_items, _i32p),
                                 _ptr(ids, _i32p), _ptr(sc, _f32p)))
        return ids, sc

    def eval_metrics(self, train_indptr, train_items, test_indptr, test_items):
        a = np.ascontiguousarray(train_indptr, np.int64); b = np.ascontiguousarray(train_items, np.int32)
        c = np.ascontiguousarray(test_indptr, np.int64); d = np.ascontiguousarray(test_items, np.int32)
        out = np.empty(6, np.float64); sums = np.empty(7, np.float64)
        _ck(lib().heat_eval_metrics(self._h, _ptr(a, _i64p), _ptr(b, _i32p), _ptr(c, _i64p), _ptr(d, _i32p),
                                    _ptr(out, _f64p), _ptr(sums, _f64p)))
        names = ["Recall(k=20)", "Recall(k=50)", "NDCG(k=20)", "NDCG(k=50)", "HitRate(k=20)", "HitRate(k=50)"]
        return dict(zip(names, out.tolist())), sums

    # -- multi-GPU
    @staticmethod
    def comm_unique_id():
        buf = np.zeros(128, np.uint8)
        _ck(lib().heat_comm_unique_id(_ptr(buf, _u8p)))
        return buf

    def comm_init(self, rank, world, uid):
        uid = np.ascontiguousarray(uid, np.uint8)
        _ck(lib().heat_comm_init(self._h, rank, world, _ptr(uid, _u8p)))

    def comm_mode(self):
        """0 no communicator, 1 NCCL all-reduce, 2 fused peer-memory exchange kernel."""
        v = C.c_int(); _ck(lib().heat_comm_mode(self._h, C.byref(v))); return v.value

    def comm_sync_items(self):
        _ck(lib().heat_comm_sync_items(self._h))

    def allreduce_f64(self, vals):
        v = np.ascontiguousarray(vals, np.float64).copy()
        _ck(lib().heat_comm_allreduce_f64(self._h, _ptr(v, _f64p), v.size))
        return v

    def set_sync_interval(self, samples):
        _ck(lib().heat_set_sync_interval(self._h, samples))

    def last_exchange(self):
        """(exchanges of the last epoch, their total device time in ms)"""
        n, ms = C.c_uint64(), C.c_float()
        _ck(lib().heat_last_exchange(self._h, C.byref(n), C.byref(ms)))
        return n.value, ms.value


class LocalGroup:
    """n contexts on one device trained as n ranks would be (heat_local_group_*): rank r = ctxs[r] holds shard r."""

    def __init__(self, ctxs, share=0):
        self.ctxs = list(ctxs)
        self._arr = (C.c_void_p * len(self.ctxs))(*[c._h for c in self.ctxs])
        _ck(lib().heat_local_group_init(self._arr, len(self.ctxs)))
        if share:       # experiment knob: bit 0 W0, bit 1 item grad rows, bit 2 item table held ONCE for all members
            _ck(lib().heat_local_group_share(self._arr, len(self.ctxs), share))

    def train_epoch(self):
        loss = C.c_double()
        _ck(lib().heat_local_group_train_epoch(self._arr, len(self.ctxs), C.byref(loss)))
        return loss.value
