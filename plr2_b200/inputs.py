"""Input step before the hot path, on the device.

Batched equivalents of what ``datasets/ycb/dataset.py:143-146,177-197,215-219,251-289`` (and ``src/loaders_v2/dataset_ycb.py:
91-119``: ``_compute_choose``, ``pointcloud``) do per frame with numpy on the host:

    mask_and_bbox      depth + label + class id -> object mask, ``get_bbox`` crop, pixel count
    choose_from_mask   mask crop -> ``choose`` (F,1,N) int64   (nonzero, wrap pad / order-preserving subset)
    pointcloud         depth + choose + intrinsics -> ``cloud`` (F,N,3) float32
    transform_target   model_points . R^T + t (+ add_t) -> ``target`` (F,M,3) float32

All go through the C-ABI (``plr2_mask_bbox``, ``plr2_choose_from_mask``, ``plr2_backproject``, ``plr2_transform_target``);
CPU tensors raise ``Plr2Error`` - there is no host fallback.
"""
import torch

from . import _cabi


def _cuda(t, what):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise _cabi.Plr2Error("inputs.%s: CUDA tensors required (no CPU fallback in plr2_b200)" % what)
    return t


def choose_from_mask(mask, bbox, num_pt, keep=None, seed=0):
    """mask (F,H,W) uint8/bool (object label AND valid depth), bbox (F,4) int (rmin,rmax,cmin,cmax) ->
    (choose (F,1,num_pt) int64, count (F,) int32).  ``keep`` (F,K) uint8: the reference's shuffled ``c_mask``
    per frame (entry k keeps the k-th non-zero pixel) for frames with more than num_pt object pixels; without
    it the subset is drawn on the device from ``seed``."""
    mask = _cuda(mask, "choose_from_mask")
    if mask.dtype == torch.bool:
        mask = mask.to(torch.uint8)
    if mask.dtype != torch.uint8 or mask.dim() != 3:
        raise ValueError("inputs.choose_from_mask: mask must be (F,H,W) uint8/bool")
    mask = mask.contiguous()
    frames, h, w = mask.shape
    dev = mask.device
    bbox = _cuda(bbox, "choose_from_mask").to(torch.int32).contiguous()
    if tuple(bbox.shape) != (frames, 4):
        raise ValueError("inputs.choose_from_mask: bbox must be (F,4)")
    choose = torch.empty(frames, 1, num_pt, device=dev, dtype=torch.int64)
    count = torch.empty(frames, device=dev, dtype=torch.int32)
    if frames == 0:
        return choose, count
    kp, kstride = None, 0
    if keep is not None:
        keep = _cuda(keep, "choose_from_mask").to(torch.uint8).contiguous()
        if keep.dim() != 2 or keep.shape[0] != frames:
            raise ValueError("inputs.choose_from_mask: keep must be (F,K)")
        if keep.shape[1] < h * w:                      # the kernel indexes keep by non-zero rank (< H*W)
            keep = torch.nn.functional.pad(keep, (0, h * w - keep.shape[1]))
        kp, kstride = _cabi.ptr(keep), keep.shape[1]
    L = _cabi.lib()
    ws_bytes = L.plr2_choose_workspace_bytes(frames, h * w)
    ws = torch.empty(ws_bytes // 4, device=dev, dtype=torch.int32)
    with torch.cuda.device(dev):
        rc = L.plr2_choose_from_mask(_cabi.ptr(mask), h, w, _cabi.ptr(bbox), frames, num_pt, kp, kstride,
                                     int(seed) & 0xFFFFFFFFFFFFFFFF, _cabi.ptr(choose), _cabi.ptr(count),
                                     _cabi.ptr(ws), ws_bytes, _cabi.stream_ptr(dev))
    _cabi.check(rc, "plr2_choose_from_mask")
    return choose, count


def pointcloud(depth, bbox, choose, cam, add_t=None):
    """depth (F,H,W) uint16 (stored as int16 bits or int32) / float32, choose (F,1,N) or (F,N) int64,
    cam (F,5) = (cx, cy, fx, fy, cam_scale), add_t (F,3) float64 or None -> cloud (F,N,3) float32."""
    depth = _cuda(depth, "pointcloud")
    frames, h, w = depth.shape
    dev = depth.device
    if depth.dtype in (torch.int16, torch.uint16):
        d, is_u16 = depth.contiguous(), 1
    elif depth.dtype in (torch.int32, torch.int64, torch.uint8):
        d, is_u16 = depth.to(torch.float32).contiguous(), 0          # exact: depth values are < 2^16
    elif depth.dtype == torch.float32:
        d, is_u16 = depth.contiguous(), 0
    else:
        raise ValueError("inputs.pointcloud: unsupported depth dtype %s" % depth.dtype)
    bbox = _cuda(bbox, "pointcloud").to(torch.int32).contiguous()
    ch = _cuda(choose, "pointcloud").to(torch.int64).reshape(frames, -1).contiguous()
    n = ch.shape[1]
    cam = _cuda(cam, "pointcloud").to(torch.float32).contiguous()
    if tuple(cam.shape) != (frames, 5) or tuple(bbox.shape) != (frames, 4):
        raise ValueError("inputs.pointcloud: cam must be (F,5) and bbox (F,4)")
    at = None
    if add_t is not None:
        at = _cuda(add_t, "pointcloud").to(torch.float64).contiguous()
        if tuple(at.shape) != (frames, 3):
            raise ValueError("inputs.pointcloud: add_t must be (F,3)")
    cloud = torch.empty(frames, n, 3, device=dev, dtype=torch.float32)
    if frames == 0:
        return cloud
    with torch.cuda.device(dev):
        rc = _cabi.lib().plr2_backproject(_cabi.ptr(d), is_u16, h, w, _cabi.ptr(bbox), _cabi.ptr(ch), _cabi.ptr(cam),
                                          _cabi.ptr(at) if at is not None else None, frames, n, _cabi.ptr(cloud),
                                          _cabi.stream_ptr(dev))
    _cabi.check(rc, "plr2_backproject")
    return cloud


def transform_target(model_points, target_r, target_t, add_t=None):
    """model_points (F,M,3) float32, target_r (F,3,3), target_t (F,3), add_t (F,3) or None (float64) ->
    target (F,M,3) float32 = model_points . R^T + t (+ add_t)."""
    mp = _cuda(model_points, "transform_target").to(torch.float32).contiguous()
    frames, m, _ = mp.shape
    dev = mp.device
    R = _cuda(target_r, "transform_target").to(torch.float64).reshape(frames, 9).contiguous()
    t = _cuda(target_t, "transform_target").to(torch.float64).reshape(frames, 3).contiguous()
    at = None
    if add_t is not None:
        at = _cuda(add_t, "transform_target").to(torch.float64).reshape(frames, 3).contiguous()
    out = torch.empty(frames, m, 3, device=dev, dtype=torch.float32)
    if frames == 0:
        return out
    with torch.cuda.device(dev):
        rc = _cabi.lib().plr2_transform_target(_cabi.ptr(mp), _cabi.ptr(R), _cabi.ptr(t),
                                               _cabi.ptr(at) if at is not None else None, frames, m,
                                               _cabi.ptr(out), _cabi.stream_ptr(dev))
    _cabi.check(rc, "plr2_transform_target")
    return out


def mask_and_bbox(depth, label, obj_id):
    """depth (F,H,W) uint16 bits (int16/uint16) or float32, label (F,H,W) uint8, obj_id (F,) int ->
    (mask (F,H,W) uint8 = (label == obj_id) & (depth != 0), bbox (F,4) int32 = get_bbox(label == obj_id),
    count (F,) int32 = mask.sum()) - datasets/ycb/dataset.py:143-146 and get_bbox :251-289."""
    depth = _cuda(depth, "mask_and_bbox")
    label = _cuda(label, "mask_and_bbox")
    if label.dtype != torch.uint8 or label.dim() != 3 or label.shape != depth.shape:
        raise ValueError("inputs.mask_and_bbox: label must be (F,H,W) uint8 with depth's shape")
    frames, h, w = depth.shape
    dev = depth.device
    if depth.dtype in (torch.int16, torch.uint16):
        d, is_u16 = depth.contiguous(), 1
    elif depth.dtype == torch.float32:
        d, is_u16 = depth.contiguous(), 0
    else:
        d, is_u16 = depth.to(torch.float32).contiguous(), 0
    oid = _cuda(obj_id, "mask_and_bbox").reshape(frames).to(torch.int32).contiguous()
    mask = torch.empty(frames, h, w, device=dev, dtype=torch.uint8)
    bbox = torch.empty(frames, 4, device=dev, dtype=torch.int32)
    count = torch.empty(frames, device=dev, dtype=torch.int32)
    if frames == 0:
        return mask, bbox, count
    with torch.cuda.device(dev):
        rc = _cabi.lib().plr2_mask_bbox(_cabi.ptr(d), is_u16, _cabi.ptr(label.contiguous()), _cabi.ptr(oid), frames, h, w,
                                        _cabi.ptr(mask), _cabi.ptr(bbox), _cabi.ptr(count), _cabi.stream_ptr(dev))
    _cabi.check(rc, "plr2_mask_bbox")
    return mask, bbox, count
