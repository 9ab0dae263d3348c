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


def _knn_plan(lib, batch, n_query, slots):
    ns, part = ctypes.c_int(0), ctypes.c_int(0)
    assert lib.plr2_knn_plan(batch, n_query, slots, ctypes.byref(ns), ctypes.byref(part)) == 0
    return ns.value, part.value


@pytest.mark.parametrize("batch,n_query", [(1, 1), (1, 127), (1, 129), (1, 2300), (512, 2300), (8, 500_000),
                                           (1, 2_300_000), (3, 1025), (65535, 7)])
def test_knn_plan_covers_every_query_once(lib, batch, n_query):
    """Host-side split of kernel (4): parts are whole lane rows (multiples of 128 queries), at most 9 queries
    per lane, they tile [0, n_query) exactly, and no split is empty."""
    for slots in (1, 148 * 4, 148 * 5):
        ns, part = _knn_plan(lib, batch, n_query, slots)
        assert ns >= 1 and part >= 128 and part % 128 == 0
        assert part // 128 <= 9
        assert (ns - 1) * part < n_query <= ns * part


def test_knn_plan_prefers_full_waves(lib):
    """C2 x 8 (500 000 queries, 8 frames) on 592 resident CTAs: one 8-query round per CTA; 512 frames of
    2300 queries: 9 queries per lane (8 + 1) in two parts, not three rounds of 6."""
    ns, part = _knn_plan(lib, 8, 500_000, 592)
    assert part == 8 * 128 and ns == 489
    ns, part = _knn_plan(lib, 512, 2300, 592)
    assert (ns, part) == (2, 9 * 128)
    # a single small frame is spread over the SMs: one query per lane
    ns, part = _knn_plan(lib, 1, 2300, 592)
    assert (ns, part) == (18, 128)
    assert lib.plr2_knn_plan(0, 10, 10, None, None) != 0


def test_training_workspace_holds_the_tensor_core_scratch(lib):
    """plr2_net_workspace_bytes(training=1) accounts for the split-bf16 operand planes and the FP32 staging area
    of the tensor-core training contractions: 3 x 1920 x (max(rows padded to 64, 1024) + 2048) floats."""
    for frames, n in ((1, 500), (8, 1000), (3, 200)):
        rows = (frames * n + 63) // 64 * 64
        scratch = 3 * 1920 * (max(rows, 1024) + 2048) * 4
        for net in (0, 1):
            train = lib.plr2_net_workspace_bytes(net, frames, n, 21, 1)
            infer_fp32_layout = lib.plr2_net_workspace_bytes(net, frames, n, 21, 0)
            assert train >= scratch
            assert train - scratch <= 2 * infer_fp32_layout + (1 << 22)


def test_loss_dropins_default_to_fast_ranking_knn_dropin_to_exact():
    from plr2_b200 import ops
    from plr2_b200.lib.knn import KNearestNeighbor
    from plr2_b200.lib.loss import Loss
    from plr2_b200.lib.loss_refiner import Loss_refine
    assert ops.KNN_LOSS_DEFAULT == ops.KNN_FAST
    assert Loss(500, [7, 8]).knn_mode == ops.KNN_FAST and Loss_refine(500, [7, 8]).knn_mode == ops.KNN_FAST
    assert Loss(500, [7, 8], ops.KNN_EXACT).knn_mode == ops.KNN_EXACT
    assert KNearestNeighbor.mode == ops.KNN_EXACT
    assert (ops.PREC_FP32, ops.PREC_BF16X2, ops.PREC_BF16X2_ALL) == (0, 1, 2)
