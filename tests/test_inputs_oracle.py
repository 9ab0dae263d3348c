"""CPU: sanity of the input-step oracle (numpy restatement of datasets/ycb/dataset.py:177-197,215-219)."""
import numpy as np

from oracle import inputs_oracle as IO


def test_choose_wrap_and_subset():
    mask = np.zeros((6, 8), dtype=np.uint8)
    mask[2, 3] = mask[3, 1] = mask[4, 6] = 1
    ch, cnt = IO.compute_choose(mask, (1, 6, 0, 8), 7)
    # crop rows 1..5, width 8: flat indices (2-1)*8+3, (3-1)*8+1, (4-1)*8+6, wrapped to length 7
    assert cnt == 3 and ch.tolist() == [11, 17, 30, 11, 17, 30, 11]
    ch, cnt = IO.compute_choose(mask, (1, 6, 0, 8), 2, c_mask=np.array([1, 0, 1]))
    assert cnt == 3 and ch.tolist() == [11, 30]


def test_pointcloud_pinhole_model_and_types():
    depth = np.full((480, 640), 10000, dtype=np.uint16)
    bbox = (100, 140, 200, 280)
    choose = np.array([0, 81, 40 * 80 - 1])
    cloud = IO.pointcloud(depth, bbox, choose, 312.9869, 241.3109, 1066.778, 1067.487, np.uint16(10000))
    assert cloud.dtype == np.float32 and cloud.shape == (3, 3)
    rows, cols = np.array([100, 101, 139]), np.array([200, 201, 279])
    np.testing.assert_allclose(cloud[:, 2], 1.0)
    np.testing.assert_allclose(cloud[:, 0], (cols - 312.9869) / 1066.778, rtol=1e-6)
    np.testing.assert_allclose(cloud[:, 1], (rows - 241.3109) / 1067.487, rtol=1e-6)


def test_transform_target_is_rigid():
    rng = np.random.default_rng(0)
    pts = rng.normal(size=(50, 3))
    R = np.linalg.qr(rng.normal(size=(3, 3)))[0]
    t = rng.normal(size=(1, 3))
    out = IO.transform_target(pts, R, t)
    np.testing.assert_allclose(out, (R @ pts.T).T + t, rtol=1e-5, atol=1e-6)
    assert out.dtype == np.float32


def test_get_bbox_snaps_to_multiples_of_40_and_stays_inside():
    lab = np.zeros((480, 640), dtype=bool)
    lab[100:141, 200:279] = True                       # 41 x 79 -> 80 x 80 around the centre
    rmin, rmax, cmin, cmax = IO.get_bbox(lab)
    assert (rmax - rmin, cmax - cmin) == (80, 80) and rmin <= 100 and rmax >= 141 and cmin <= 200 and cmax >= 279
    lab[:] = False
    lab[0:10, 630:640] = True                          # corner: shifted back inside the image
    rmin, rmax, cmin, cmax = IO.get_bbox(lab)
    assert (rmin, rmax, cmin, cmax) == (0, 40, 600, 640)
    lab[:] = False
    lab[100:140, 200:280] = True                       # already on the borders: unchanged
    assert IO.get_bbox(lab) == (100, 140, 200, 280)
