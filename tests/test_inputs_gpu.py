"""GPU: input-step kernels (choose from mask, depth back-projection, target transform) through the C-ABI
against the numpy restatement of datasets/ycb/dataset.py:177-197,215-219.  Index work is bit-exact, the
float32 back-projection is bit-exact, the float64 target transform is within one float32 ulp."""
import numpy as np
import pytest
import torch

from oracle import inputs_oracle as IO
from plr2_b200 import _cabi, inputs

pytestmark = pytest.mark.gpu

H, W = 480, 640
CAM = (312.9869, 241.3109, 1066.778, 1067.487)          # cx, cy, fx, fy of datasets/ycb/dataset.py:65-68


def blob_mask(rng, h, w, fill):
    """A connected-ish object mask with holes (invalid depth)."""
    m = np.zeros((H, W), dtype=np.uint8)
    r0, c0 = int(rng.integers(0, H - h)), int(rng.integers(0, W - w))
    m[r0:r0 + h, c0:c0 + w] = rng.random((h, w)) < fill
    return m, (r0, c0)


def make_frames(rng, specs):
    masks, bboxes = [], []
    for (h, w, fill, pad) in specs:
        m, (r0, c0) = blob_mask(rng, h, w, fill)
        rmin, rmax = max(0, r0 - pad), min(H, r0 + h + pad)
        cmin, cmax = max(0, c0 - pad), min(W, c0 + w + pad)
        masks.append(m); bboxes.append((rmin, rmax, cmin, cmax))
    return np.stack(masks), np.array(bboxes, dtype=np.int32)


SPECS = [(120, 160, 0.7, 20), (20, 30, 0.5, 10), (3, 3, 1.0, 40), (200, 200, 0.9, 0), (60, 90, 0.02, 5), (40, 40, 0.0, 4)]


@pytest.mark.parametrize("num_pt", [1000, 500, 7])
def test_choose_with_reference_c_mask_is_bit_exact(num_pt):
    rng = np.random.default_rng(num_pt)
    masks, bboxes = make_frames(rng, SPECS)
    frames = len(SPECS)
    keep = np.zeros((frames, H * W), dtype=np.uint8)
    ref, counts = [], []
    for f in range(frames):
        cnt = int(masks[f][bboxes[f][0]:bboxes[f][1], bboxes[f][2]:bboxes[f][3]].sum())
        c_mask = None
        if cnt > num_pt:                                      # dataset.py:179-181
            c_mask = np.zeros(cnt, dtype=int); c_mask[:num_pt] = 1; rng.shuffle(c_mask)
            keep[f, :cnt] = c_mask
        if cnt == 0:
            ref.append(np.zeros(num_pt, dtype=np.int64))      # the reference raises (np.pad of an empty array)
        else:
            ref.append(IO.compute_choose(masks[f], bboxes[f], num_pt, c_mask)[0])
        counts.append(cnt)
    choose, count = inputs.choose_from_mask(torch.from_numpy(masks).cuda(), torch.from_numpy(bboxes).cuda(), num_pt,
                                            keep=torch.from_numpy(keep).cuda())
    assert tuple(choose.shape) == (frames, 1, num_pt) and choose.dtype == torch.int64
    np.testing.assert_array_equal(count.cpu().numpy(), counts)
    np.testing.assert_array_equal(choose[:, 0].cpu().numpy(), np.stack(ref))


def test_choose_device_random_subset_properties():
    rng = np.random.default_rng(1)
    masks, bboxes = make_frames(rng, [(120, 160, 0.7, 20), (200, 200, 0.9, 0), (20, 30, 0.5, 10)])
    num_pt = 1000
    m, b = torch.from_numpy(masks).cuda(), torch.from_numpy(bboxes).cuda()
    c1, cnt = inputs.choose_from_mask(m, b, num_pt, seed=123)
    c1b, _ = inputs.choose_from_mask(m, b, num_pt, seed=123)
    c2, _ = inputs.choose_from_mask(m, b, num_pt, seed=124)
    assert torch.equal(c1, c1b)                                # deterministic in the seed
    c1n, c2n = c1[:, 0].cpu().numpy(), c2[:, 0].cpu().numpy()
    for f in range(2):                                         # frames with more than num_pt pixels
        full = masks[f][bboxes[f][0]:bboxes[f][1], bboxes[f][2]:bboxes[f][3]].flatten().nonzero()[0]
        assert cnt[f].item() == len(full) > num_pt
        assert np.all(np.diff(c1n[f]) > 0)                     # ascending, no repeats: exactly num_pt distinct
        assert np.isin(c1n[f], full).all()                     # a subset of the object pixels
        assert not np.array_equal(c1n[f], c2n[f])              # a different seed gives a different subset
        # spread over the whole list, not a prefix: the mean rank is near the middle
        ranks = np.searchsorted(full, c1n[f])
        assert 0.4 * len(full) < ranks.mean() < 0.6 * len(full)
    full = masks[2][bboxes[2][0]:bboxes[2][1], bboxes[2][2]:bboxes[2][3]].flatten().nonzero()[0]
    np.testing.assert_array_equal(c1n[2], np.pad(full, (0, num_pt - len(full)), 'wrap'))   # few pixels: wrap


@pytest.mark.parametrize("dtype", ["u16", "f32"])
@pytest.mark.parametrize("noise", [False, True])
def test_pointcloud_bit_exact(dtype, noise):
    rng = np.random.default_rng(7)
    masks, bboxes = make_frames(rng, SPECS[:5])
    frames, num_pt = 5, 1000
    depth = rng.integers(3000, 20000, size=(frames, H, W)).astype(np.uint16)
    cam_scale = np.uint16(10000)                               # meta['factor_depth'][0][0]
    add_t = rng.uniform(-0.03, 0.03, size=(frames, 3)) if noise else None
    chooses, ref = [], []
    for f in range(frames):
        cnt = int(masks[f][bboxes[f][0]:bboxes[f][1], bboxes[f][2]:bboxes[f][3]].sum())
        c_mask = None
        if cnt > num_pt:
            c_mask = np.zeros(cnt, dtype=int); c_mask[:num_pt] = 1; rng.shuffle(c_mask)
        ch, _ = IO.compute_choose(masks[f], bboxes[f], num_pt, c_mask)
        chooses.append(ch)
        d = depth[f] if dtype == "u16" else depth[f].astype(np.float32)
        ref.append(IO.pointcloud(d, bboxes[f], ch, *CAM, cam_scale, None if add_t is None else add_t[f]))
    cam = np.array([[*CAM, 10000.0]] * frames, dtype=np.float32)
    if dtype == "u16":
        dt = torch.from_numpy(depth.view(np.int16)).cuda()      # same bits; the kernel reads uint16
    else:
        dt = torch.from_numpy(depth.astype(np.float32)).cuda()
    cloud = inputs.pointcloud(dt, torch.from_numpy(bboxes[:frames]).cuda(), torch.from_numpy(np.stack(chooses)).cuda()[:, None],
                              torch.from_numpy(cam).cuda(), None if add_t is None else torch.from_numpy(add_t).cuda())
    got, want = cloud.cpu().numpy(), np.stack(ref)
    assert got.dtype == np.float32 and got.shape == want.shape
    np.testing.assert_array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("noise", [False, True])
def test_transform_target(noise):
    rng = np.random.default_rng(9)
    frames, m = 4, 2600
    model = ((rng.random((frames, m, 3)) - 0.5) * 0.3).astype(np.float32)
    R = np.stack([np.linalg.qr(rng.normal(size=(3, 3)))[0] for _ in range(frames)])
    t = rng.normal(size=(frames, 3)) * 0.5
    add_t = rng.uniform(-0.03, 0.03, size=(frames, 3)) if noise else None
    want = np.stack([IO.transform_target(model[f].astype(np.float64), R[f], t[f][None], None if add_t is None else add_t[f])
                     for f in range(frames)])
    got = inputs.transform_target(torch.from_numpy(model).cuda(), torch.from_numpy(R).cuda(), torch.from_numpy(t).cuda(),
                                  None if add_t is None else torch.from_numpy(add_t).cuda()).cpu().numpy()
    # float64 arithmetic rounded once: identical up to the summation order of np.dot (<= 1 float32 ulp)
    ulp = np.abs(got.view(np.int32).astype(np.int64) - want.view(np.int32).astype(np.int64))
    assert ulp.max() <= 1 and (ulp == 0).mean() > 0.999


def test_feeds_the_network_inputs():
    """choose/cloud produced on the device have the shapes and dtypes PoseNet.forward takes
    (lib/network.py:95: x (bs,N,3), choose (bs,1,N) int64)."""
    rng = np.random.default_rng(2)
    masks, bboxes = make_frames(rng, [(120, 160, 0.7, 20)] * 3)
    m, b = torch.from_numpy(masks).cuda(), torch.from_numpy(bboxes).cuda()
    choose, _ = inputs.choose_from_mask(m, b, 1000, seed=5)
    depth = torch.from_numpy(rng.integers(3000, 20000, size=(3, H, W)).astype(np.float32)).cuda()
    cam = torch.tensor([[*CAM, 10000.0]] * 3, device="cuda")
    cloud = inputs.pointcloud(depth, b, choose, cam)
    assert tuple(cloud.shape) == (3, 1000, 3) and cloud.dtype == torch.float32 and torch.isfinite(cloud).all()
    crop = (b[:, 1] - b[:, 0]) * (b[:, 3] - b[:, 2])
    assert (choose[:, 0].max(dim=1).values < crop).all() and (choose >= 0).all()


def test_no_cpu_fallback():
    with pytest.raises(_cabi.Plr2Error):
        inputs.choose_from_mask(torch.zeros(1, 4, 4, dtype=torch.uint8), torch.zeros(1, 4, dtype=torch.int32), 4)
    with pytest.raises(_cabi.Plr2Error):
        inputs.pointcloud(torch.zeros(1, 4, 4), torch.zeros(1, 4), torch.zeros(1, 1, 4), torch.zeros(1, 5))


def test_mask_and_bbox_bit_exact():
    """mask = (label == obj) & (depth != 0) and get_bbox of the label mask against the numpy restatement of
    dataset.py:143-145,251-289: blobs at the image borders, on exact multiples of 40, single pixels, absent object."""
    rng = np.random.default_rng(21)
    cases = [(0, 0, 37, 55), (470, 630, 10, 10), (100, 200, 40, 80), (100, 200, 41, 79), (0, 600, 480, 40), (200, 300, 1, 1),
             (5, 5, 300, 500), (430, 0, 50, 640), (17, 333, 121, 161)]
    frames = len(cases) + 1
    label = rng.integers(0, 5, size=(frames, H, W)).astype(np.uint8) * 0            # background 0
    depth = rng.integers(0, 3, size=(frames, H, W)).astype(np.uint16) * 4000        # a third of the pixels invalid
    obj = np.arange(3, 3 + frames).astype(np.int32)
    for f, (r0, c0, h, w) in enumerate(cases):
        blob = rng.random((h, w)) < 0.8
        blob[0, :] |= True; blob[-1, :] |= True; blob[:, 0] |= True; blob[:, -1] |= True     # tight extent known
        label[f, r0:r0 + h, c0:c0 + w][blob] = obj[f]
        label[f, (r0 + 200) % H, (c0 + 300) % W] = obj[f] + 1 if f % 2 else 0            # another class elsewhere
    # last frame: object absent
    mask, bbox, count = inputs.mask_and_bbox(torch.from_numpy(depth.view(np.int16)).cuda(), torch.from_numpy(label).cuda(),
                                             torch.from_numpy(obj).cuda())
    mask, bbox, count = mask.cpu().numpy(), bbox.cpu().numpy(), count.cpu().numpy()
    for f in range(frames - 1):
        m, ml = IO.object_mask(depth[f], label[f], obj[f])
        np.testing.assert_array_equal(mask[f], m.astype(np.uint8))
        assert count[f] == int(m.sum())
        assert tuple(bbox[f]) == tuple(int(v) for v in IO.get_bbox(ml)), (f, bbox[f], IO.get_bbox(ml))
        rmin, rmax, cmin, cmax = bbox[f]
        assert 0 <= rmin < rmax <= H and 0 <= cmin < cmax <= W and (rmax - rmin) % 40 == 0 and (cmax - cmin) % 40 == 0
    assert tuple(bbox[-1]) == (0, 0, 0, 0) and count[-1] == 0 and mask[-1].sum() == 0
    # float32 depth gives the same
    mask2, bbox2, _ = inputs.mask_and_bbox(torch.from_numpy(depth.astype(np.float32)).cuda(), torch.from_numpy(label).cuda(),
                                           torch.from_numpy(obj).cuda())
    assert np.array_equal(mask2.cpu().numpy(), mask) and np.array_equal(bbox2.cpu().numpy(), bbox)


def test_input_chain_on_device():
    """depth + label -> mask/bbox -> choose -> cloud, all on the device, equals the numpy chain frame by frame."""
    rng = np.random.default_rng(31)
    frames, num_pt = 4, 1000
    label = np.zeros((frames, H, W), dtype=np.uint8)
    depth = rng.integers(3000, 20000, size=(frames, H, W)).astype(np.uint16)
    depth[rng.random((frames, H, W)) < 0.1] = 0
    obj = np.array([7, 2, 15, 9], dtype=np.int32)
    for f, (r0, c0, h, w) in enumerate([(50, 60, 130, 170), (300, 400, 60, 45), (0, 0, 25, 25), (200, 100, 200, 333)]):
        label[f, r0:r0 + h, c0:c0 + w][rng.random((h, w)) < 0.85] = obj[f]
    dd, dl, do = torch.from_numpy(depth.view(np.int16)).cuda(), torch.from_numpy(label).cuda(), torch.from_numpy(obj).cuda()
    mask, bbox, count = inputs.mask_and_bbox(dd, dl, do)
    keep = np.zeros((frames, H * W), dtype=np.uint8)
    ref_choose, ref_cloud = [], []
    for f in range(frames):
        m, ml = IO.object_mask(depth[f], label[f], obj[f])
        bb = IO.get_bbox(ml)
        cnt = int(m[bb[0]:bb[1], bb[2]:bb[3]].sum())
        c_mask = None
        if cnt > num_pt:
            c_mask = np.zeros(cnt, dtype=int); c_mask[:num_pt] = 1; rng.shuffle(c_mask); keep[f, :cnt] = c_mask
        ch, _ = IO.compute_choose(m, bb, num_pt, c_mask)
        ref_choose.append(ch)
        ref_cloud.append(IO.pointcloud(depth[f], bb, ch, *CAM, np.uint16(10000)))
    choose, cnt = inputs.choose_from_mask(mask, bbox, num_pt, keep=torch.from_numpy(keep).cuda())
    cam = torch.tensor([[*CAM, 10000.0]] * frames, device="cuda")
    cloud = inputs.pointcloud(dd, bbox, choose, cam)
    np.testing.assert_array_equal(choose[:, 0].cpu().numpy(), np.stack(ref_choose))
    np.testing.assert_array_equal(cloud.cpu().numpy().view(np.uint32), np.stack(ref_cloud).view(np.uint32))
