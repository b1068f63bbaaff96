"""CPU: the C-ABI library loads, exports every symbol include/bvc_gpu.h declares, the struct layouts agree,
and every entry point fails loudly (no CPU fallback) when there is no GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import basevarc_b200 as B
from basevarc_b200 import _abi, caller

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "bvc_gpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bvc_[a-z0-9_]+)\s*\(", src)))


def test_exports_match_header():
    lib = B.load_library()
    names = declared_functions()
    assert set(names) == set(caller.EXPORTS)
    for n in names:
        assert getattr(lib, n) is not None


def test_struct_layouts():
    assert _abi.SITE_RESULT_DTYPE.itemsize == 176
    assert C.sizeof(_abi.Tile) == 8 + 7 * 8 + 2 * 8 + 8        # + packed8, qual_dict, n_qual_dict + reserved (ABI 2)
    assert C.sizeof(_abi.Params) == 8 + 8 * 3 + 8 + 8 + 8 + 8
    lib = B.load_library()
    p = _abi.Params()
    lib.bvc_default_params(C.byref(p), 1000, 0.001)
    q = B.default_params(1000, 0.001)
    assert (p.abi_version, p.n_samples, p.min_af, p.lrt_threshold, p.em_epsilon, p.em_max_iter) == (
        q.abi_version, q.n_samples, q.min_af, q.lrt_threshold, q.em_epsilon, q.em_max_iter)
    lib.bvc_default_params(C.byref(p), 1000000, 0.01)
    assert p.min_af == 100.0 / 1000000


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(B.BvcError, match="no CUDA device"):
        B.BaseTypeCaller(100)
    with pytest.raises(B.BvcError):
        B.measure_fp64_peak(0)


def test_make_code_roundtrip():
    b, s, i = np.meshgrid(np.arange(8), np.arange(2), np.arange(2), indexing="ij")
    c = B.make_code(b, s, i)
    assert ((c & 7) == b).all() and (((c >> 3) & 1) == s).all() and (((c >> 4) & 1) == i).all() and c.max() < 32
