"""Shared helpers for the parity tests (oracle = checker, never the thing tested)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TRAIN_CASES = ["train_d64_k16", "train_d128_k16", "train_d64_k16_dups", "train_d32_k8"]
# large-negative regime (train_kernel_bigk.cuh); their `negs` table holds the first 72 samples only
BIGK_CASES = ["train_d128_k1000", "train_d64_k200", "train_d32_k48"]


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: z[k] for k in z.files}


def rel_err(a, b):
    """max |a-b| relative to the LARGEST reference magnitude: a max-norm ratio, not an element-wise relative error
    (near-zero elements of a table would make the latter meaningless); the north-star's "1e-5 relative" is read in
    this norm."""
    denom = max(float(np.abs(b).max()), 1e-12)
    return float(np.abs(a.astype(np.float64) - b.astype(np.float64)).max()) / denom


def fresh_tables(g):
    return g["init_user_w"].copy(), g["init_item_w"].copy(), g["init_w0"].copy()
