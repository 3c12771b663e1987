"""CPU-side checks of the drop-in boundary: the library loads and exports every symbol
include/hg_b200.h declares; argument validation works without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def hg():
    import hectorgrapher_b200 as hg
    if not os.path.exists(hg.capi.LIB_PATH):
        hg.build()
    return hg


def test_header_symbols_are_exported(hg):
    header = open(os.path.join(ROOT, "include", "hg_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(hg_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 35
    L = C.CDLL(hg.capi.LIB_PATH)
    missing = [s for s in sorted(declared) if not hasattr(L, s)]
    assert not missing, missing
    # the ctypes table covers exactly the declared symbols
    assert declared == set(hg.capi.SYMBOLS.keys())


def test_version_and_error_strings(hg):
    L = hg.lib()
    assert b"sm_100a" in L.hg_version()
    assert isinstance(L.hg_last_error(), bytes)


def test_argument_validation_needs_no_gpu(hg):
    L = hg.lib()
    out = C.c_void_p()
    rc = L.hg_grid_upload_prob(-1.0, 0, 0, None, None, C.byref(out))
    assert rc == 1 and b"hg_grid_upload_prob" in L.hg_last_error()
    rc = L.hg_batch_create(99, 1, None, 3, None, None, None, None, 1, C.byref(out))
    assert rc == 1
    rc = L.hg_stack_build(None, 7, 3, C.byref(out))
    assert rc == 1


def test_no_cpu_fallback(hg):
    """Without a CUDA device every compute entry point must fail loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    xyz = np.zeros((1, 3), np.int32)
    v = np.array([1000], np.uint16)
    with pytest.raises(hg.HgError):
        hg.Grid.prob(0.1, xyz, v)


def test_structs_match_header_layout(hg):
    assert C.sizeof(hg.capi.FcsmOptions) == 64
    assert C.sizeof(hg.capi.NodeData) == 80
    assert C.sizeof(hg.capi.MatchResult) == 152
