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


def object_mask(rng, height, width):
    """Synthetic object mask inside its crop, shaped like the reference's `mask = mask_label * mask_depth` cropped to
    `get_bbox` (datasets/ycb/dataset.py:143-146,177): a compact blob that touches the four sides of the crop
    (the crop IS the blob's bounding box, rounded up to a multiple of 40) with ~3 % invalid-depth holes."""
    yy, xx = np.mgrid[0:height, 0:width]
    cy, cx = (height - 1) / 2.0, (width - 1) / 2.0
    ang = np.arctan2(yy - cy, xx - cx)
    # superellipse radius modulated by a few low harmonics: a convex-ish silhouette filling 60-75 % of the box
    k = rng.uniform(0.05, 0.15, 3)
    ph = rng.uniform(0, 2 * np.pi, 3)
    mod = 1.0 + sum(k[i] * np.cos((i + 2) * ang + ph[i]) for i in range(3))
    ry, rx = 0.47 * height * mod, 0.47 * width * mod
    inside = ((yy - cy) / ry) ** 4 + ((xx - cx) / rx) ** 4 <= 1.0
    holes = rng.random((height, width)) < 0.03
    return inside & ~holes


def choose_from_mask(rng, mask, n):
    """datasets/ycb/dataset.py:177-184: indices of the non-zero mask pixels of the crop, flattened row-major; more than
    `n` -> a random subset of n keeping ascending order, fewer -> np.pad(..., 'wrap')."""
    choose = mask.flatten().nonzero()[0]
    if len(choose) > n:
        c_mask = np.zeros(len(choose), dtype=int)
        c_mask[:n] = 1
        rng.shuffle(c_mask)
        choose = choose[c_mask.nonzero()]
    elif len(choose) > 0:
        choose = np.pad(choose, (0, n - len(choose)), 'wrap')
    else:
        choose = np.zeros(n, dtype=np.int64)
    return choose


def make_frame(shape: Shape, seed: int, obj=None, with_image=True, choose_mode="uniform"):
    """One frame as a dict of numpy arrays in the reference's per-frame layout (batch dim of 1
    omitted): points (N,3), choose (N,) i64, out_img (32,H,W), target (M,3), model_points (M,3),
    idx () i64, plus gt_R (3,3), gt_t (3,), gt_q (4,).  choose_mode: "uniform" = sorted random subset of the crop
    (what the golden fixtures were generated with), "mask" = pixels of a compact object blob (separate
    random streams, so every other array of the frame is identical in both modes)."""
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
    if choose_mode == "mask":      # what the reference's loader produces: the object's pixels (bench.py workload)
        choose = choose_from_mask(_rng(seed + 0x5eed), object_mask(_rng(seed + 0x3a5c), shape.height, shape.width), n)
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
