"""Seeded synthetic YCB-Video-shaped raw frames (colour, depth, label, meta) shared by
tests/golden/make_inputs_golden.py (which writes them to disk as the PNG / .mat files the reference's loader reads)
and by the tests (which regenerate the same arrays instead of storing 1.2 MB of depth + label per frame).
PNG is lossless for uint8 / uint16, so both sides see identical arrays."""
import numpy as np

H, W = 480, 640
# (frame name in the reference's file lists, seed, blobs = [(class id 1-based, row0, col0, height, width, fill)])
FRAMES = [
    ("data/0048/000001", 11, [(1, 60, 80, 150, 190, 0.93), (13, 300, 400, 70, 55, 0.85), (20, 250, 60, 90, 120, 0.9)]),
    ("data/0048/000036", 12, [(16, 0, 0, 45, 62, 0.95), (5, 200, 280, 231, 317, 0.97)]),            # blob on the image corner
    ("data/0048/000047", 13, [(21, 420, 560, 60, 80, 0.9), (9, 100, 100, 12, 9, 1.0)]),            # bottom-right corner; 108-pixel object
    ("data/0048/000083", 14, [(19, 120, 200, 40, 80, 0.88), (3, 239, 40, 161, 119, 0.8)]),          # extents on / next to multiples of 40
]


def raw_frame(seed, blobs):
    rng = np.random.Generator(np.random.PCG64(seed))
    color = rng.integers(0, 256, size=(H, W, 3), dtype=np.uint8)
    yy, xx = np.mgrid[0:H, 0:W]
    depth = (9000 + 2500 * np.sin(xx / 97.0 + seed) * np.cos(yy / 61.0) + rng.integers(-40, 41, size=(H, W))).astype(np.uint16)
    depth[rng.random((H, W)) < 0.06] = 0                           # invalid depth
    label = np.zeros((H, W), dtype=np.uint8)
    poses = np.zeros((3, 4, len(blobs)))
    for i, (cid, r0, c0, h, w, fill) in enumerate(blobs):
        sub = label[r0:r0 + h, c0:c0 + w]
        sub[rng.random(sub.shape) < fill] = cid
        q = rng.standard_normal(4)
        q /= np.linalg.norm(q)
        a, b, c, d = q
        poses[:, :3, i] = np.array([[1 - 2 * (c * c + d * d), 2 * b * c - 2 * a * d, 2 * a * c + 2 * b * d],
                                    [2 * b * c + 2 * d * a, 1 - 2 * (b * b + d * d), -2 * a * b + 2 * c * d],
                                    [-2 * a * c + 2 * b * d, 2 * a * b + 2 * c * d, 1 - 2 * (b * b + c * c)]])
        poses[:, 3, i] = [rng.uniform(-0.2, 0.2), rng.uniform(-0.2, 0.2), rng.uniform(0.6, 1.2)]
    cls = np.array([[b[0]] for b in blobs], dtype=np.uint8)
    return {"color": color, "depth": depth, "label": label, "poses": poses, "cls_indexes": cls,
            "factor_depth": np.array([[10000]], dtype=np.uint16)}
