"""CPU: the C-ABI library builds, loads and exports every symbol include/qpmpc.h declares; struct layouts agree.
No compute call is made (no GPU here) except the loud-failure check of qpmpc_create."""
import ctypes as C
import os
import re

import pytest

from quadruped_pympc_tamols_b200 import _cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    _cabi.build()
    return _cabi.lib()


def test_exports_every_declared_symbol(lib):
    hdr = open(os.path.join(ROOT, "include", "qpmpc.h")).read()
    declared = set(re.findall(r"\b(qpmpc_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_cabi.EXPORTS), declared ^ set(_cabi.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name


def test_exports_every_declared_tamols_symbol(lib):
    from quadruped_pympc_tamols_b200 import _tamols_cabi as tc
    hdr = open(os.path.join(ROOT, "include", "tamols.h")).read()
    declared = set(re.findall(r"\b(tamols_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(tc.EXPORTS), declared ^ set(tc.EXPORTS)
    L = tc.lib()
    for name in declared:
        assert hasattr(L, name), name
    assert C.sizeof(tc.TamolsParams) == 8 * 24
    assert C.sizeof(tc.TamolsLegJob) == 24 + 8 * 19
    assert C.sizeof(tc.TamolsResult) == 8 + 8 * 10


def test_tamols_create_fails_loudly_without_gpu(lib):
    import torch
    from quadruped_pympc_tamols_b200 import _tamols_cabi as tc
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(tc.TamolsError, match="no CUDA device|CPU fallback"):
        tc.Handle(tc.make_params({}, "go2"))


def test_struct_sizes_match_header():
    assert C.sizeof(_cabi.QpmpcResult) == 4 * 52
    assert C.sizeof(_cabi.QpmpcBatch) == 8 + 14 * 8
    assert C.sizeof(_cabi.QpmpcConfig) == 4 * (8 + 1 + 9 + 3 + 32 + 12 + 2 + 3 + 1 + 1 + 8 + 2 + 2)


def test_version_string(lib):
    assert b"sm_100a" in lib.qpmpc_version()


def test_create_fails_loudly_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    cfg = _cabi.make_config(mass=10.0, inertia=[[1, 0, 0], [0, 1, 0], [0, 0, 1]])
    with pytest.raises(_cabi.QpmpcError, match="no CUDA device|CPU fallback"):
        _cabi.Handle(cfg)


def test_bad_config_rejected():
    with pytest.raises(_cabi.QpmpcError):
        _cabi.make_config(mass=1.0, inertia=[[1, 0, 0], [0, 1, 0], [0, 0, 1]], sampling_method="nope")
    with pytest.raises(_cabi.QpmpcError):
        _cabi.make_config(mass=1.0, inertia=[[1, 0, 0], [0, 1, 0], [0, 0, 1]], horizon=64)
