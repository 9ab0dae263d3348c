"""Seeded synthetic YCB / LineMOD-shaped inputs and weights.

numpy's PCG64 streams are stable across platforms, so the same seed gives the same frame in the
golden-fixture generator (run where /root/reference exists), in the tests and in bench.py.
Shapes follow datasets/ycb/dataset.py:227-232 and datasets/linemod/dataset.py of the reference.
"""
from dataclasses import dataclass

import numpy as np

YCB_SYM = [12, 15, 18, 19, 20]      # datasets/ycb/dataset.py:89
LINEMOD_SYM = [7, 8]                # datasets/linemod/dataset.py:88


@dataclass(frozen=True)
class Shape:
    num_points: int
    num_obj: int
    num_mesh: int
    height: int
    width: int
    sym_list: tuple


YCB = Shape(1000, 21, 500, 120, 120, tuple(YCB_SYM))
YCB_REFINE = Shape(1000, 21, 2300, 120, 120, tuple(YCB_SYM))      # num_pt_mesh_large
LINEMOD = Shape(500, 13, 500, 80, 80, tuple(LINEMOD_SYM))


def _rng(seed):
    return np.random.Generator(np.random.PCG64(int(seed)))


def random_rotation(rng):
    q = rng.standard_normal(4)
    q /= np.linalg.norm(q)
    return quat_to_rot_np(q), q


def quat_to_rot_np(q):
    w, x, y, z = [float(v) for v in q]
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * x * y - 2 * w * z, 2 * w * y + 2 * x * z],
        [2 * x * y + 2 * z * w, 1 - 2 * (x * x + z * z), -2 * w * x + 2 * y * z],
        [-2 * w * y + 2 * x * z, 2 * w * x + 2 * y * z, 1 - 2 * (x * x + y * y)]], dtype=np.float64)


def mesh_points(rng, m):
    """Points on a surface of revolution (bowl/cylinder-like), r in [0.04,0.08] m, h in [0.05,0.15] m."""
    r0 = rng.uniform(0.04, 0.08)
    h = rng.uniform(0.05, 0.15)
    z = rng.uniform(-0.5, 0.5, m)
    phi = rng.uniform(0, 2 * np.pi, m)
    r = r0 * (0.75 + 0.25 * np.cos(np.pi * z))      # gently bulging profile
    return np.stack([r * np.cos(phi), r * np.sin(phi), h * z], 1).astype(np.float32)


def make_frame(shape: Shape, seed: int, obj=None, with_image=True):
    """One frame as a dict of numpy arrays in the reference's per-frame layout (batch dim of 1
    omitted): points (N,3), choose (N,) i64, out_img (32,H,W), target (M,3), model_points (M,3),
    idx () i64, plus gt_R (3,3), gt_t (3,), gt_q (4,)."""
    rng = _rng(seed)
    n, m = shape.num_points, shape.num_mesh
    hw = shape.height * shape.width
    model = mesh_points(rng, m)
    rot, q = random_rotation(rng)
    t = np.array([rng.uniform(-0.2, 0.2), rng.uniform(-0.2, 0.2), rng.uniform(0.6, 1.2)])
    target = (model.astype(np.float64) @ rot.T + t).astype(np.float32)
    pick = rng.integers(0, m, n)
    points = (target[pick] + rng.normal(0, 0.002, (n, 3))).astype(np.float32)
    if hw >= n:
        choose = np.sort(rng.choice(hw, n, replace=False))
    else:  # wrap-pad like datasets/ycb/dataset.py:178-184
        choose = np.sort(np.resize(rng.permutation(hw), n))
    frame = {
        "points": points, "choose": choose.astype(np.int64), "target": target, "model_points": model,
        "idx": np.array(rng.integers(0, shape.num_obj) if obj is None else obj, dtype=np.int64),
        "gt_R": rot.astype(np.float32), "gt_t": t.astype(np.float32), "gt_q": q.astype(np.float32),
    }
    if with_image:
        logits = rng.standard_normal((32, shape.height, shape.width)).astype(np.float32)
        mx = logits.max(0, keepdims=True)
        lse = mx + np.log(np.exp(logits - mx).sum(0, keepdims=True))
        frame["out_img"] = (logits - lse).astype(np.float32)     # log_softmax over channels
    return frame


def make_predictions(frame, seed, noise_r=0.1, noise_t=0.01):
    """Loss-only inputs: pred_r = q_gt + 0.1 randn, pred_t = (t_gt - points) + 0.01 randn,
    pred_c = sigmoid(randn)  (SURVEY.md §8d)."""
    rng = _rng(seed)
    n = frame["points"].shape[0]
    pred_r = (frame["gt_q"][None, :] + noise_r * rng.standard_normal((n, 4))).astype(np.float32)
    pred_t = ((frame["gt_t"][None, :] - frame["points"]) + noise_t * rng.standard_normal((n, 3))).astype(np.float32)
    pred_c = (1.0 / (1.0 + np.exp(-rng.standard_normal((n, 1))))).astype(np.float32)
    return pred_r, pred_t, pred_c


# parameter (name, shape) lists in the reference's state-dict naming (lib/network.py:39-52,70-93,136-186)
def posenet_param_shapes(num_obj):
    s = [("feat.conv1", 64, 3), ("feat.conv2", 128, 64), ("feat.e_conv1", 64, 32), ("feat.e_conv2", 128, 64),
         ("feat.conv5", 512, 256), ("feat.conv6", 1024, 512)]
    for h in "rtc":
        s.append((f"conv1_{h}", 640, 1408))
    for h in "rtc":
        s.append((f"conv2_{h}", 256, 640))
    for h in "rtc":
        s.append((f"conv3_{h}", 128, 256))
    s += [("conv4_r", num_obj * 4, 128), ("conv4_t", num_obj * 3, 128), ("conv4_c", num_obj, 128)]
    out = []
    for name, o, i in s:
        out.append((name + ".weight", (o, i, 1)))
        out.append((name + ".bias", (o,)))
    return out


def refiner_param_shapes(num_obj):
    conv = [("feat.conv1", 64, 3), ("feat.conv2", 128, 64), ("feat.e_conv1", 64, 32), ("feat.e_conv2", 128, 64),
            ("feat.conv5", 512, 384), ("feat.conv6", 1024, 512)]
    lin = [("conv1_r", 512, 1024), ("conv1_t", 512, 1024), ("conv2_r", 128, 512), ("conv2_t", 128, 512),
           ("conv3_r", num_obj * 4, 128), ("conv3_t", num_obj * 3, 128)]
    out = []
    for name, o, i in conv:
        out.append((name + ".weight", (o, i, 1)))
        out.append((name + ".bias", (o,)))
    for name, o, i in lin:
        out.append((name + ".weight", (o, i)))
        out.append((name + ".bias", (o,)))
    return out


def make_params(kind, num_obj, seed):
    """Weights with torch's default Conv1d/Linear init distribution, U(-1/sqrt(fan_in), +1/sqrt(fan_in)),
    drawn from a numpy stream (so they do not depend on module construction order)."""
    rng = _rng(seed)
    shapes = posenet_param_shapes(num_obj) if kind == "posenet" else refiner_param_shapes(num_obj)
    params = {}
    fan_in = None
    for name, shp in shapes:
        if name.endswith(".weight"):
            fan_in = shp[1]
        bound = 1.0 / np.sqrt(fan_in)
        params[name] = rng.uniform(-bound, bound, shp).astype(np.float32)
    return params
