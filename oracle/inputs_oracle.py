"""ORACLE — TEST INFRASTRUCTURE ONLY (never imported by plr2_b200/).

numpy restatement of the reference's per-frame input step, statement for statement:

    compute_choose    <- datasets/ycb/dataset.py:177-184  (= src/loaders_v2/dataset_ycb.py:91-106)
    pointcloud        <- datasets/ycb/dataset.py:186-196  (= src/loaders_v2/dataset_ycb.py:108-119)
    transform_target  <- datasets/ycb/dataset.py:215-219,229

PINNED: tests/golden/make_inputs_golden.py runs the reference's OWN loader (PoseDataset.__getitem__ and get_bbox,
imported from /root/reference) on seeded synthetic YCB-Video files and commits its outputs
(tests/golden/inputs_ycb.npz); tests/test_inputs_golden.py checks these functions against them exactly (box, choose
and the float32 cloud bit for bit).  The functions are the loader's numpy calls on explicit arrays (numpy itself is
the third-party arithmetic, so the float32/float64 promotion rules are the reference's own).  xmap holds the row index and ymap the column index
(dataset.py:83-84).
"""
import numpy as np


def compute_choose(mask, bbox, num_pt, c_mask=None):
    """mask (H,W) bool/uint8, bbox (rmin,rmax,cmin,cmax); c_mask: the shuffled 0/1 vector of the reference
    (drawn by the caller so that the test controls the RNG).  Returns (choose (num_pt,), count)."""
    rmin, rmax, cmin, cmax = [int(v) for v in bbox]
    choose = mask[rmin:rmax, cmin:cmax].flatten().nonzero()[0]
    count = len(choose)
    if count > num_pt:
        assert c_mask is not None and len(c_mask) >= count and int(np.sum(c_mask[:count] != 0)) == num_pt
        choose = choose[np.asarray(c_mask[:count]).nonzero()]
    else:
        choose = np.pad(choose, (0, num_pt - count), 'wrap')
    return choose, count


def pointcloud(depth, bbox, choose, cam_cx, cam_cy, cam_fx, cam_fy, cam_scale, add_t=None):
    rmin, rmax, cmin, cmax = [int(v) for v in bbox]
    h, w = depth.shape
    xmap = np.array([[j for i in range(w)] for j in range(h)])
    ymap = np.array([[i for i in range(w)] for j in range(h)])
    depth_masked = depth[rmin:rmax, cmin:cmax].flatten()[choose][:, np.newaxis].astype(np.float32)
    xmap_masked = xmap[rmin:rmax, cmin:cmax].flatten()[choose][:, np.newaxis].astype(np.float32)
    ymap_masked = ymap[rmin:rmax, cmin:cmax].flatten()[choose][:, np.newaxis].astype(np.float32)
    pt2 = depth_masked / cam_scale
    pt0 = (ymap_masked - cam_cx) * pt2 / cam_fx
    pt1 = (xmap_masked - cam_cy) * pt2 / cam_fy
    cloud = np.concatenate((pt0, pt1, pt2), axis=1)
    if add_t is not None:
        cloud = np.add(cloud, add_t)
    return cloud.astype(np.float32)


def transform_target(model_points, target_r, target_t, add_t=None):
    target = np.dot(model_points, target_r.T)
    if add_t is not None:
        target = np.add(target, target_t + add_t)
    else:
        target = np.add(target, target_t)
    return target.astype(np.float32)


# datasets/ycb/dataset.py:247-289, statement for statement
border_list = [-1, 40, 80, 120, 160, 200, 240, 280, 320, 360, 400, 440, 480, 520, 560, 600, 640, 680]


def get_bbox(label, img_width=480, img_length=640):
    rows = np.any(label, axis=1)
    cols = np.any(label, axis=0)
    rmin, rmax = np.where(rows)[0][[0, -1]]
    cmin, cmax = np.where(cols)[0][[0, -1]]
    rmax += 1
    cmax += 1
    r_b = rmax - rmin
    for tt in range(len(border_list)):
        if r_b > border_list[tt] and r_b < border_list[tt + 1]:
            r_b = border_list[tt + 1]
            break
    c_b = cmax - cmin
    for tt in range(len(border_list)):
        if c_b > border_list[tt] and c_b < border_list[tt + 1]:
            c_b = border_list[tt + 1]
            break
    center = [int((rmin + rmax) / 2), int((cmin + cmax) / 2)]
    rmin = center[0] - int(r_b / 2)
    rmax = center[0] + int(r_b / 2)
    cmin = center[1] - int(c_b / 2)
    cmax = center[1] + int(c_b / 2)
    if rmin < 0:
        delt = -rmin
        rmin = 0
        rmax += delt
    if cmin < 0:
        delt = -cmin
        cmin = 0
        cmax += delt
    if rmax > img_width:
        delt = rmax - img_width
        rmax = img_width
        rmin -= delt
    if cmax > img_length:
        delt = cmax - img_length
        cmax = img_length
        cmin -= delt
    return rmin, rmax, cmin, cmax


def object_mask(depth, label, obj_id):
    """dataset.py:143-145."""
    import numpy.ma as ma
    mask_depth = ma.getmaskarray(ma.masked_not_equal(depth, 0))
    mask_label = ma.getmaskarray(ma.masked_equal(label, obj_id))
    return mask_label * mask_depth, mask_label
