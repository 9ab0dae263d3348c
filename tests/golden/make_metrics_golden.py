"""Generates tests/golden/metrics_add_adds.npz from the REFERENCE's own ADD / ADDS
(src/estimation/errors.py:53-84), imported from /root/reference/src.  Run in the build container only
(the GPU box has no /root/reference):  python tests/golden/make_metrics_golden.py
"""
import os
import sys
import types

import numpy as np

try:
    import cv2  # noqa: F401  (errors.py imports it for rotation_error only)
except Exception:                                   # pragma: no cover
    sys.modules["cv2"] = types.ModuleType("cv2")
sys.path.insert(0, "/root/reference/src")
from estimation.errors import ADD, ADDS, rotation_error, translation_error   # noqa: E402
from estimation.state import State_R3xQuat, State_SE3, points       # noqa: E402
from scipy.spatial.transform import Rotation         # noqa: E402


def random_pose(rng, angle_scale, t_scale):
    T = np.eye(4)
    T[:3, :3] = Rotation.from_rotvec(rng.normal(size=3) * angle_scale).as_matrix()
    T[:3, 3] = rng.normal(size=3) * t_scale
    return T


def main():
    rng = np.random.default_rng(20260)
    out = {}
    for name, m, frames, ang, ts in (("small", 64, 6, 0.3, 0.05), ("ycb", 500, 4, 0.05, 0.01), ("far", 97, 3, 2.0, 0.5)):
        model = (rng.random((frames, m, 3)) - 0.5) * 0.2
        p1 = np.stack([random_pose(rng, ang, ts) for _ in range(frames)])
        p2 = np.stack([random_pose(rng, ang, ts) for _ in range(frames)])
        add = np.array([ADD(points(model[f]), State_SE3(p1[f]), State_SE3(p2[f])) for f in range(frames)])
        adds = np.array([ADDS(points(model[f]), State_SE3(p1[f]), State_SE3(p2[f])) for f in range(frames)])
        out.update({f"{name}_model": model, f"{name}_pose1": p1, f"{name}_pose2": p2,
                    f"{name}_add": add, f"{name}_adds": adds})
    # rotation / translation errors (errors.py:12-51) on State_R3xQuat poses, incl. identical, tiny and ~180 deg cases
    n = 24
    quats = rng.normal(size=(n, 4)); quats /= np.linalg.norm(quats, axis=1, keepdims=True)
    quats2 = rng.normal(size=(n, 4)); quats2 /= np.linalg.norm(quats2, axis=1, keepdims=True)
    quats2[0] = quats[0]                                                    # identical
    quats2[1] = quats[1] + 1e-4 * rng.normal(size=4); quats2[1] /= np.linalg.norm(quats2[1])   # tiny angle
    quats2[2] = np.array([quats[2][3], -quats[2][2], quats[2][1], -quats[2][0]])                # orthogonal quaternion: 180 deg
    ts, ts2 = rng.normal(size=(n, 3)), rng.normal(size=(n, 3))
    ts2[0] = ts[0]
    pa = [State_R3xQuat(list(ts[i]) + list(quats[i])) for i in range(n)]
    pb = [State_R3xQuat(list(ts2[i]) + list(quats2[i])) for i in range(n)]
    out["err_pose1"] = np.stack([State_SE3(p).state for p in pa])
    out["err_pose2"] = np.stack([State_SE3(p).state for p in pb])
    out["err_rot"] = np.array([rotation_error(pa[i], pb[i]) for i in range(n)])
    out["err_trans"] = np.array([translation_error(pa[i], pb[i]) for i in range(n)])
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "metrics_add_adds.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
