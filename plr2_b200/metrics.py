"""Evaluation metrics after the hot path, computed on the device.

Mirrors the reference's evaluation helpers:
  * ``ADD`` / ``ADDS``  - ``src/estimation/errors.py:53-67`` / ``:69-84`` (model points under two poses;
    ADD-S = mean distance of every ``pose2`` point to its nearest ``pose1`` point);
  * ``pose_errors``     - ``src/estimation/errors.py:12-51`` (rotation angle and translation distance between poses);
  * ``auc``             - ``replace_ycb_toolbox/plot_accuracy_keyframe.m:38-52`` with ``VOCap`` ``:150-170``
    (AUC of the accuracy-threshold curve up to 0.1 m and the "< 2 cm" rate).

Poses are ``(F, 12)`` tensors, row-major ``R`` followed by ``t`` - the layout ``DenseFusionPipeline`` returns
(``last_poses`` / ``infer(..., return_pose=True)``) - or ``(F, 4, 4)`` / ``(F, 3, 4)`` homogeneous matrices.
Everything runs through the C-ABI (``plr2_pose_metric``, ``plr2_auc_f32``); there is no CPU fallback.
"""
import torch

from . import _cabi


def _pose12(p, frames_hint=None):
    if not isinstance(p, torch.Tensor) or not p.is_cuda:
        raise _cabi.Plr2Error("metrics: poses must be CUDA tensors (no CPU fallback in plr2_b200)")
    p = p.float()
    if p.dim() == 1:
        p = p.unsqueeze(0)
    if p.dim() == 3 and p.shape[-1] == 4 and p.shape[-2] in (3, 4):     # (F, 3|4, 4) homogeneous
        p = torch.cat([p[:, :3, :3].reshape(-1, 9), p[:, :3, 3]], dim=1)
    if p.dim() != 2 or p.shape[1] != 12:
        raise ValueError("metrics: pose must be (F,12), (F,3,4) or (F,4,4); got %s" % (tuple(p.shape),))
    return p.contiguous()


def pose_metrics(model_points, pose1, pose2, want_add=True, want_adds=True):
    """Per-frame ``(add, adds)`` (each ``(F,)`` float32 or ``None``) for model_points ``(M,3)`` (shared by all
    frames) or ``(F,M,3)``."""
    a, b = _pose12(pose1), _pose12(pose2)
    if a.shape != b.shape:
        raise ValueError("metrics: pose1/pose2 shapes differ: %s vs %s" % (tuple(a.shape), tuple(b.shape)))
    if not model_points.is_cuda:
        raise _cabi.Plr2Error("metrics: model_points must be a CUDA tensor (no CPU fallback in plr2_b200)")
    mp = model_points.float().contiguous()
    frames = a.shape[0]
    if mp.dim() == 2:
        stride, m = 0, mp.shape[0]
    elif mp.dim() == 3 and mp.shape[0] == frames:
        stride, m = mp.shape[1] * 3, mp.shape[1]
    else:
        raise ValueError("metrics: model_points must be (M,3) or (F,M,3); got %s" % (tuple(mp.shape),))
    if mp.shape[-1] != 3:
        raise ValueError("metrics: model_points last dimension must be 3")
    dev = a.device
    add = torch.empty(frames, device=dev, dtype=torch.float32) if want_add else None
    adds = torch.empty(frames, device=dev, dtype=torch.float32) if want_adds else None
    if frames == 0:
        return add, adds
    L = _cabi.lib()
    ws_bytes = L.plr2_pose_metric_workspace_bytes(frames)
    ws = torch.empty(ws_bytes, device=dev, dtype=torch.uint8)
    with torch.cuda.device(dev):
        rc = L.plr2_pose_metric(_cabi.ptr(mp), stride, _cabi.ptr(a), _cabi.ptr(b), frames, m,
                                _cabi.ptr(add) if want_add else None, _cabi.ptr(adds) if want_adds else None,
                                _cabi.ptr(ws), ws_bytes, _cabi.stream_ptr(dev))
    _cabi.check(rc, "plr2_pose_metric")
    return add, adds


def ADD(model_points, pose1, pose2):
    """``src/estimation/errors.py:53-67`` for F frames at once."""
    return pose_metrics(model_points, pose1, pose2, True, False)[0]


def ADDS(model_points, pose1, pose2):
    """``src/estimation/errors.py:69-84`` for F frames at once."""
    return pose_metrics(model_points, pose1, pose2, False, True)[1]


def pose_errors(pose1, pose2):
    """``(rotation_error, translation_error)`` per frame, each ``(F,)`` float32: ``src/estimation/errors.py:12-32``
    (norm of the Rodrigues vector of ``R1 R2^T``, i.e. the rotation angle in radians) and ``:34-51``."""
    a, b = _pose12(pose1), _pose12(pose2)
    if a.shape != b.shape:
        raise ValueError("metrics: pose1/pose2 shapes differ: %s vs %s" % (tuple(a.shape), tuple(b.shape)))
    frames, dev = a.shape[0], a.device
    rot = torch.empty(frames, device=dev, dtype=torch.float32)
    tr = torch.empty(frames, device=dev, dtype=torch.float32)
    if frames == 0:
        return rot, tr
    with torch.cuda.device(dev):
        rc = _cabi.lib().plr2_pose_errors(_cabi.ptr(a), _cabi.ptr(b), frames, _cabi.ptr(rot), _cabi.ptr(tr),
                                          _cabi.stream_ptr(dev))
    _cabi.check(rc, "plr2_pose_errors")
    return rot, tr


def auc(distances, max_distance=0.1, below=0.02):
    """``(AUC, rate_below)`` as a ``(2,)`` float64 CUDA tensor (no host sync): the toolbox's ``VOCap(d, accuracy)``
    and ``c / n`` of ``plot_accuracy_keyframe.m:38-52`` (which prints both times 100)."""
    if not isinstance(distances, torch.Tensor) or not distances.is_cuda:
        raise _cabi.Plr2Error("metrics.auc: distances must be a CUDA tensor (no CPU fallback in plr2_b200)")
    d = distances.float().contiguous().view(-1)
    dev = d.device
    out = torch.empty(2, device=dev, dtype=torch.float64)
    L = _cabi.lib()
    n = d.numel()
    ws_bytes = L.plr2_auc_workspace_bytes(n)
    ws = torch.empty((ws_bytes + 7) // 8, device=dev, dtype=torch.float64)
    with torch.cuda.device(dev):
        rc = L.plr2_auc_f32(_cabi.ptr(d) if n else None, n, float(max_distance), float(below), _cabi.ptr(out),
                            _cabi.ptr(ws), ws_bytes, _cabi.stream_ptr(dev))
    _cabi.check(rc, "plr2_auc_f32")
    return out
