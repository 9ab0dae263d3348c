"""Row f-2 pinned on the reference's OWN loader: tests/golden/inputs_ycb.npz holds what datasets/ycb/dataset.py
(PoseDataset.__getitem__ :97-232, get_bbox :251-289) returned for seeded synthetic YCB-Video files
(tests/golden/make_inputs_golden.py).  CPU: the numpy oracle reproduces it exactly.  GPU: the device kernels
(plr2_mask_bbox, plr2_choose_from_mask, plr2_backproject, plr2_transform_target) reproduce it through the C-ABI -
indices and the float32 cloud bit for bit, the float64 target transform within one float32 ulp."""
import os
import sys

import numpy as np
import pytest
import torch

from oracle import inputs_oracle as IO

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import inputs_synth as S  # noqa: E402

G = dict(np.load(os.path.join(HERE, "golden", "inputs_ycb.npz")))
CAM = (312.9869, 241.3109, 1066.778, 1067.487)          # cam_cx_1 .. cam_fy_1, datasets/ycb/dataset.py:65-68 (folder 0048 < 60)
CASES = sorted({k.rsplit("_", 1)[0] for k in G if k.endswith("_cloud")})


def _case(key):
    k = int(key[1:key.index("_")])
    name, seed, blobs = S.FRAMES[k]
    fr = S.raw_frame(seed, blobs)
    g = {n: G[f"{key}_{n}"] for n in ("cloud", "choose", "target", "model_points", "idx", "bbox", "num_pt", "pose_col")}
    return fr, g


def _kept(mask, bbox, choose, num_pt):
    """The reference's shuffled c_mask (dataset.py:178-181), recovered from the pixels it kept."""
    rmin, rmax, cmin, cmax = [int(v) for v in bbox]
    all_idx = mask[rmin:rmax, cmin:cmax].flatten().nonzero()[0]
    if len(all_idx) <= num_pt:
        return None, len(all_idx)
    c_mask = np.isin(all_idx, choose).astype(int)
    assert c_mask.sum() == num_pt
    return c_mask, len(all_idx)


def test_fixture_covers_the_interesting_cases():
    assert len(CASES) == 16
    counts, boxes = [], set()
    for key in CASES:
        fr, g = _case(key)
        m, _ = IO.object_mask(fr["depth"], fr["label"], int(g["idx"]) + 1)
        counts.append(int(m.sum()))
        boxes.add(tuple(int(v) for v in g["bbox"]))
    assert min(counts) < 500 and max(counts) > 1000           # wrap-padded and subsampled `choose`
    assert any(b[0] == 0 and b[2] == 0 for b in boxes) and any(b[1] == 480 and b[3] == 640 for b in boxes)   # boxes pushed back inside


@pytest.mark.parametrize("key", CASES)
def test_oracle_reproduces_the_reference_loader(key):
    fr, g = _case(key)
    obj_id, num_pt = int(g["idx"]) + 1, int(g["num_pt"])
    mask, mask_label = IO.object_mask(fr["depth"], fr["label"], obj_id)
    bbox = IO.get_bbox(mask_label)
    assert tuple(int(v) for v in bbox) == tuple(int(v) for v in g["bbox"])
    c_mask, _ = _kept(mask, bbox, g["choose"][0], num_pt)
    choose, _ = IO.compute_choose(mask, bbox, num_pt, c_mask)
    assert np.array_equal(choose, g["choose"][0])
    cloud = IO.pointcloud(fr["depth"], bbox, choose, *CAM, fr["factor_depth"][0][0])
    assert np.array_equal(cloud.view(np.uint32), g["cloud"].view(np.uint32))
    col = int(g["pose_col"])
    # the loader hands float64 model points to np.dot (dataset.py:215) and returns their float32 image: replay from those
    target = IO.transform_target(g["model_points"].astype(np.float64), fr["poses"][:, :3, col],
                                 np.array([fr["poses"][:, 3, col]]))
    assert np.allclose(target, g["target"], rtol=0, atol=2e-7)


@pytest.mark.gpu
def test_device_input_step_reproduces_the_reference_loader():
    from plr2_b200 import inputs
    frs, gs = zip(*[_case(k) for k in CASES])
    F = len(CASES)
    for num_pt in sorted({int(g["num_pt"]) for g in gs}):
        sel = [i for i in range(F) if int(gs[i]["num_pt"]) == num_pt]
        depth = np.stack([frs[i]["depth"] for i in sel])
        label = np.stack([frs[i]["label"] for i in sel])
        obj = np.array([int(gs[i]["idx"]) + 1 for i in sel], dtype=np.int32)
        dd = torch.from_numpy(depth.view(np.int16)).cuda()
        mask, bbox, count = inputs.mask_and_bbox(dd, torch.from_numpy(label).cuda(), torch.from_numpy(obj).cuda())
        np.testing.assert_array_equal(bbox.cpu().numpy(), np.stack([gs[i]["bbox"] for i in sel]))
        keep = np.zeros((len(sel), S.H * S.W), dtype=np.uint8)
        for j, i in enumerate(sel):
            m, _ = IO.object_mask(frs[i]["depth"], frs[i]["label"], int(obj[j]))
            np.testing.assert_array_equal(mask[j].cpu().numpy(), m.astype(np.uint8))
            c_mask, cnt = _kept(m, gs[i]["bbox"], gs[i]["choose"][0], num_pt)
            if c_mask is not None:
                keep[j, :cnt] = c_mask
        choose, cnt = inputs.choose_from_mask(mask, bbox, num_pt, keep=torch.from_numpy(keep).cuda())
        np.testing.assert_array_equal(choose[:, 0].cpu().numpy(), np.stack([gs[i]["choose"][0] for i in sel]))
        cam = torch.tensor([[*CAM, 10000.0]] * len(sel), device="cuda")
        cloud = inputs.pointcloud(dd, bbox, choose, cam)
        np.testing.assert_array_equal(cloud.cpu().numpy().view(np.uint32), np.stack([gs[i]["cloud"] for i in sel]).view(np.uint32))
        for j, i in enumerate(sel):                 # model sets differ in size (500 / 2600): one call per frame
            col = int(gs[i]["pose_col"])
            R = torch.from_numpy(frs[i]["poses"][:, :3, col].copy())[None].cuda()
            t = torch.from_numpy(frs[i]["poses"][:, 3, col].copy())[None].cuda()
            tgt = inputs.transform_target(torch.from_numpy(gs[i]["model_points"])[None].cuda(), R, t)[0].cpu().numpy()
            assert np.allclose(tgt, gs[i]["target"], rtol=0, atol=2e-7)
