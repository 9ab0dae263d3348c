"""CPU: the C-ABI library loads and exports every symbol include/plr2_b200.h declares; the ctypes
table binds exactly that set; argument validation paths that need no GPU behave."""
import ctypes
import os
import re

import pytest

from plr2_b200 import _cabi, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "plr2_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(plr2_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return _cabi.lib()


def test_header_symbols_exported_and_bound(lib):
    syms = declared_symbols()
    assert len(syms) >= 9
    raw = ctypes.CDLL(_cabi.LIB_PATH)
    for s in syms:
        assert hasattr(raw, s), f"{s} declared in include/plr2_b200.h but not exported"
    assert sorted(_cabi.SIGNATURES) == syms, "ctypes SIGNATURES out of sync with the header"


def test_version_and_error_string(lib):
    assert lib.plr2_abi_version() == 1
    assert isinstance(lib.plr2_last_error_string(), bytes)


def test_argument_validation_without_gpu(lib):
    # shape errors are detected before any CUDA call
    assert lib.plr2_knn_f32(None, None, None, None, 1, 3, 0, 10, 1, 0, None, 0, None) == -2
    assert b"bad shape" in lib.plr2_last_error_string()
    assert lib.plr2_knn_f32(None, None, None, None, 1, 3, 5, 10, 1, 0, None, 0, None) == -1   # null pointers
    assert lib.plr2_knn_f32(None, None, None, None, 1, 3, 5, 0, 1, 0, None, 0, None) == 0     # empty query set
    assert lib.plr2_knn_f32(None, None, None, None, 0, 3, 5, 9, 1, 0, None, 0, None) == 0     # empty batch
    assert lib.plr2_pose_loss_fwd(*([None] * 8), 21, 1, 0, 500, 0.015, 0, 0, *([None] * 5)) == -2
    assert lib.plr2_refine_loss_fwd(*([None] * 7), 21, 1, 1000, 9000, 0, *([None] * 6), None, 0, None) in (-1, -3)
    assert lib.plr2_refine_loss_workspace_bytes(64, 2300) == 128 * 2 + 64 * 24 * 13 * 4


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_cabi, "_lib", None)
    monkeypatch.setattr(_cabi, "LIB_PATH", "/nonexistent/libplr2_b200.so")
    with pytest.raises(_cabi.Plr2Error):
        _cabi.lib()


def test_cpu_tensors_are_rejected():
    import torch
    from plr2_b200.lib.knn import knn_indices
    with pytest.raises(_cabi.Plr2Error):
        knn_indices(torch.zeros(1, 3, 4), torch.zeros(1, 3, 5))


def test_grad_regions_describe_the_head_weights(lib):
    """plr2_net_grad_regions (no GPU needed): the early range is inside the packed gradient, starts after the trunk
    weights and covers most of PoseNet's parameters (what the overlapped all-reduce sends first)."""
    import ctypes as C
    for net, num_obj in ((0, 21), (1, 21), (0, 13)):
        a, b = C.c_size_t(0), C.c_size_t(0)
        assert lib.plr2_net_grad_regions(net, num_obj, C.byref(a), C.byref(b)) == 0
        total = lib.plr2_net_packed_floats(net, num_obj) // 2
        assert 0 < a.value < b.value <= total
        if net == 0:
            assert (b.value - a.value) / total > 0.8
    assert lib.plr2_net_grad_regions(0, 21, None, None) != 0
