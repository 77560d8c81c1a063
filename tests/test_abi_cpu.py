"""CPU checks of the C-ABI library: it loads, exports every declared symbol, its host sampler twin
agrees with the oracle, and it refuses to run without a GPU (no fallback)."""
import os
import re

import numpy as np
import pytest

import oracle
from heat_b200 import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "heat_b200.h")).read()
    declared = set(re.findall(r"\b(heat_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"heat_ctx", "heat_config"}
    assert declared == set(abi.SYMBOLS), declared ^ set(abi.SYMBOLS)
    L = abi.lib()
    for s in declared:
        assert hasattr(L, s), s
    assert L.heat_abi_version() == 2


def test_host_sampler_twin_matches_oracle():
    rng = np.random.default_rng(0)
    for _ in range(200):
        seed, ep, s = int(rng.integers(1, 2**62)), int(rng.integers(0, 100)), int(rng.integers(0, 2**40))
        k, n = int(rng.integers(0, 64)), int(rng.integers(1, 2**31 - 1))
        r = oracle.philox4x32_10([s & 0xffffffff, s >> 32, ep, k >> 2], [seed & 0xffffffff, seed >> 32])[k & 3]
        assert abi.philox_draw(seed, ep, s, k, n) == (r * n) >> 32


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(abi.HeatError, match="no CPU fallback"):
        abi.Context(64, 16, 10, 10)


def test_host_tile_sampler_twin_matches_oracle():
    rng = np.random.default_rng(1)
    clicks = np.stack([np.arange(300) // 6, rng.integers(0, 500, 300)], 1).astype(np.uint64)
    for tile, refresh in ((16, 20), (512, 8192), (1, 1)):
        t = oracle.neg_table(11, 3, 12345, clicks, 8, 500, tile_size=tile, refresh_interval=refresh)
        tw = np.array([[abi.tile_draw(11, 3, 12345 + s, k, 500, tile, refresh) for k in range(8)] for s in range(300)], dtype=np.uint64)
        clash = tw == clicks[:, 1:2]
        assert np.array_equal(t[~clash], tw[~clash])
    assert abi.tile_draw(11, 3, 99, 5, 500, 0, 1) == abi.philox_draw(11, 3, 99, 5, 500)


def test_ctypes_config_layout_matches_the_c_header(tmp_path):
    """heat_b200/abi.py mirrors `heat_config` by hand: field offsets and the struct size must be the C compiler's."""
    import ctypes
    import subprocess
    fields = [f[0] for f in abi.HeatConfig._fields_]
    src = '#include <stdio.h>\n#include <stddef.h>\n#include "heat_b200.h"\nint main(void) {\n'
    for f in fields:
        src += f'  printf("{f} %zu\\n", offsetof(heat_config, {f}));\n'
    src += '  printf("sizeof %zu\\n", sizeof(heat_config));\n  return 0;\n}\n'
    c = tmp_path / "layout.c"
    c.write_text(src)
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(c), "-o", str(exe)])
    out = dict(line.split() for line in subprocess.check_output([str(exe)], text=True).splitlines())
    for f in fields:
        assert int(out[f]) == getattr(abi.HeatConfig, f).offset, f
    assert int(out["sizeof"]) == ctypes.sizeof(abi.HeatConfig)
    # every field of the header is mirrored (a field added to the header only would shift nothing but be unreachable)
    hdr = open(os.path.join(ROOT, "include", "heat_b200.h")).read()
    body = hdr[hdr.index("typedef struct heat_config {"):hdr.index("} heat_config;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    declared = re.findall(r"\b(?:uint64_t|uint32_t|float|const uint64_t \*)\s*\**(\w+);", body)
    assert declared == fields, (declared, fields)
