"""ORACLE — TEST INFRASTRUCTURE ONLY (never imported by plr2_b200/).

numpy float64 restatement of the reference's evaluation metrics:

    add, adds   <- src/estimation/errors.py:53-67, 69-84 (points.apply_transform: src/estimation/state.py:235-247)
    rotation_error, translation_error <- src/estimation/errors.py:12-51
    vocap, auc  <- replace_ycb_toolbox/plot_accuracy_keyframe.m:38-52 (sort, accuracy, "< 2 cm") and
                   VOCap :150-170

Pinning: add/adds/rotation_error/translation_error are checked against outputs of the reference's own
``estimation.errors`` functions (imported
from /root/reference/src by tests/golden/make_metrics_golden.py, fixture tests/golden/metrics_add_adds.npz)
and against its known-answer cases in src/estimation/main.py:117-131.  The AUC script is MATLAB and cannot be
run here: PARITY UNPINNED for ``auc`` beyond hand-computed known answers (tests/test_metrics_oracle.py).
"""
import numpy as np


def _apply(pose, pts):
    """pose (4,4) or (3,4) homogeneous, pts (M,3) -> (M,3): R p + t."""
    pose = np.asarray(pose, dtype=np.float64)
    pts = np.asarray(pts, dtype=np.float64)
    return pts @ pose[:3, :3].T + pose[:3, 3]


def add(model_points, pose1, pose2):
    m1, m2 = _apply(pose1, model_points), _apply(pose2, model_points)
    return float(np.mean(np.linalg.norm(m1 - m2, axis=1)))


def adds(model_points, pose1, pose2):
    """Mean over the pose2 points of the distance to the nearest pose1 point (errors.py:80-83: the tree is
    fitted on model_1 and queried with model_2)."""
    m1, m2 = _apply(pose1, model_points), _apply(pose2, model_points)
    d2 = ((m2[:, None, :] - m1[None, :, :]) ** 2).sum(-1)
    return float(np.mean(np.sqrt(d2.min(axis=1))))


def vocap(rec, prec, max_distance=0.1):
    """VOCap (:150-170): rec sorted distances (inf = miss), prec the accuracies at those ranks."""
    rec = np.asarray(rec, dtype=np.float64)
    prec = np.asarray(prec, dtype=np.float64)
    keep = np.isfinite(rec)
    rec, prec = rec[keep], prec[keep]
    if rec.size == 0:
        return 0.0            # MATLAB would fail on prec(end); the device kernel reports 0
    mrec = np.concatenate([[0.0], rec, [max_distance]])
    mpre = np.concatenate([[0.0], prec, [prec[-1]]])
    for i in range(1, mpre.size):
        mpre[i] = max(mpre[i], mpre[i - 1])
    i = np.nonzero(mrec[1:] != mrec[:-1])[0] + 1
    return float(np.sum((mrec[i] - mrec[i - 1]) * mpre[i]) * (1.0 / max_distance))


def auc(distances, max_distance=0.1, below=0.02):
    """(AUC, fraction below `below`) as plot_accuracy_keyframe.m:38-52 computes them for one class/method."""
    d = np.asarray(distances, dtype=np.float64).reshape(-1).copy()
    n = d.size
    if n == 0:
        return 0.0, 0.0
    d[~(d <= max_distance)] = np.inf          # D(D > max_distance) = inf; NaN is dropped by isfinite later
    d = np.sort(d)
    c = int(np.sum(d < below))
    accuracy = np.cumsum(np.ones(n)) / n
    return vocap(d, accuracy, max_distance), c / n


def rotation_error(pose1, pose2):
    """errors.py:12-32 on homogeneous matrices: norm of the Rodrigues vector of R1 R2^T = its rotation angle."""
    R = np.asarray(pose1, dtype=np.float64)[:3, :3] @ np.asarray(pose2, dtype=np.float64)[:3, :3].T
    v = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    return float(np.arctan2(np.linalg.norm(v), np.trace(R) - 1.0))


def translation_error(pose1, pose2):
    """errors.py:34-51."""
    return float(np.linalg.norm(np.asarray(pose1, dtype=np.float64)[:3, 3] - np.asarray(pose2, dtype=np.float64)[:3, 3]))
