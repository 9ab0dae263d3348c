"""ORACLE — TEST INFRASTRUCTURE ONLY (never imported by plr2_b200/).

Plain torch-fp32 CPU restatement of the reference's hot path, written as functions over explicit
parameter dicts (autograd supplies the reference gradients):

    posenet_feat / posenet_heads / posenet_forward   <- lib/network.py:53-68, 95-132
    refiner_forward                                  <- lib/network.py:151-168, 187-206
    quat_to_rot, pose_loss                           <- lib/loss.py:12-69
    refine_loss                                      <- lib/loss_refiner.py:12-62
    nearest_index                                    <- lib/knn (via oracle/knn_oracle.c)

Pinned against the reference's own modules (imported from /root/reference by
tests/golden/make_golden.py) through the committed fixtures in tests/golden/*.npz; see
tests/test_oracle_golden.py.  The dense contractions themselves are third-party arithmetic
(torch matmul); nothing in the reference's own tests pins them.
"""
import numpy as np
import torch

from . import knn_oracle as _knn


# ------------------------------------------------------------------------------------------------
# network
def _pw(x, w, b):
    """1x1 Conv1d on channel-major (C, N) activations: weight (out, in, 1) or (out, in)."""
    w2 = w.reshape(w.shape[0], -1)
    return w2 @ x + b[:, None]


def gather_embedding(out_img, choose):
    """emb[c, n] = out_img[c, choose[n]] over the flattened H*W map (lib/network.py:98-102)."""
    c = out_img.shape[0]
    return out_img.reshape(c, -1)[:, choose.reshape(-1)]


def posenet_feat(p, x, emb, prefix="feat."):
    """x (3, N), emb (32, N) -> (1408, N) = [pointfeat_1 | pointfeat_2 | broadcast mean of conv6]."""
    r = torch.relu
    x1 = r(_pw(x, p[prefix + "conv1.weight"], p[prefix + "conv1.bias"]))
    e1 = r(_pw(emb, p[prefix + "e_conv1.weight"], p[prefix + "e_conv1.bias"]))
    x2 = r(_pw(x1, p[prefix + "conv2.weight"], p[prefix + "conv2.bias"]))
    e2 = r(_pw(e1, p[prefix + "e_conv2.weight"], p[prefix + "e_conv2.bias"]))
    pf1 = torch.cat([x1, e1], 0)
    pf2 = torch.cat([x2, e2], 0)
    x5 = r(_pw(pf2, p[prefix + "conv5.weight"], p[prefix + "conv5.bias"]))
    x6 = r(_pw(x5, p[prefix + "conv6.weight"], p[prefix + "conv6.bias"]))
    n = x.shape[1]
    ap = x6.mean(dim=1, keepdim=True).expand(-1, n)  # AvgPool1d(num_points) + repeat
    return torch.cat([pf1, pf2, ap], 0)


def posenet_heads(p, feat, obj, num_obj):
    """(1408, N) -> pred_r (N, 4), pred_t (N, 3), pred_c (N, 1) for class `obj`."""
    r = torch.relu
    outs = {}
    for h, width in (("r", 4), ("t", 3), ("c", 1)):
        y = feat
        for layer in (1, 2, 3):
            y = r(_pw(y, p[f"conv{layer}_{h}.weight"], p[f"conv{layer}_{h}.bias"]))
        y = _pw(y, p[f"conv4_{h}.weight"], p[f"conv4_{h}.bias"])
        if h == "c":
            y = torch.sigmoid(y)
        n = y.shape[1]
        outs[h] = y.reshape(num_obj, width, n)[int(obj)].t().contiguous()
    return outs["r"], outs["t"], outs["c"]


def posenet_forward(p, out_img, x, choose, obj, num_obj):
    """One frame.  out_img (32, H, W) is the backbone output (the backbone itself is out of scope),
    x (N, 3) points, choose (N,) int64, obj int.  Returns pred_r, pred_t, pred_c, emb (32, N)."""
    emb = gather_embedding(out_img, choose)
    feat = posenet_feat(p, x.t(), emb)
    pr, pt, pc = posenet_heads(p, feat, obj, num_obj)
    return pr, pt, pc, emb.detach()


def refiner_forward(p, x, emb, obj, num_obj):
    """x (N, 3), emb (32, N) -> pred_r (1, 4), pred_t (1, 3)   (lib/network.py:151-206)."""
    r = torch.relu
    xt = x.t()
    x1 = r(_pw(xt, p["feat.conv1.weight"], p["feat.conv1.bias"]))
    e1 = r(_pw(emb, p["feat.e_conv1.weight"], p["feat.e_conv1.bias"]))
    x2 = r(_pw(x1, p["feat.conv2.weight"], p["feat.conv2.bias"]))
    e2 = r(_pw(e1, p["feat.e_conv2.weight"], p["feat.e_conv2.bias"]))
    pf3 = torch.cat([x1, e1, x2, e2], 0)  # 384 channels
    x5 = r(_pw(pf3, p["feat.conv5.weight"], p["feat.conv5.bias"]))
    x6 = r(_pw(x5, p["feat.conv6.weight"], p["feat.conv6.bias"]))
    ap = x6.mean(dim=1)  # (1024,)
    out = []
    for h, width in (("r", 4), ("t", 3)):
        y = r(p[f"conv1_{h}.weight"] @ ap + p[f"conv1_{h}.bias"])
        y = r(p[f"conv2_{h}.weight"] @ y + p[f"conv2_{h}.bias"])
        y = p[f"conv3_{h}.weight"] @ y + p[f"conv3_{h}.bias"]
        out.append(y.reshape(num_obj, width)[int(obj)].reshape(1, width))
    return out[0], out[1]


# ------------------------------------------------------------------------------------------------
# losses
def quat_to_rot(q):
    """(..., 4) normalised (w, x, y, z) -> (..., 3, 3), element expressions of lib/loss.py:17-25."""
    w, x, y, z = q.unbind(-1)
    rows = [1.0 - 2.0 * (y * y + z * z), 2.0 * x * y - 2.0 * w * z, 2.0 * w * y + 2.0 * x * z,
            2.0 * x * y + 2.0 * z * w, 1.0 - 2.0 * (x * x + z * z), -2.0 * w * x + 2.0 * y * z,
            -2.0 * w * y + 2.0 * x * z, 2.0 * w * x + 2.0 * y * z, 1.0 - 2.0 * (x * x + y * y)]
    return torch.stack(rows, -1).reshape(q.shape[:-1] + (3, 3))


def nearest_index(ref_pts, query_pts, contract=True):
    """0-based index of the nearest ref (lowest index on ties) for each query; (.., 3) tensors."""
    idx, _ = _knn.knn_oracle(ref_pts.detach().numpy().T, query_pts.detach().numpy().T, 1, contract)
    return torch.from_numpy(idx[0] - 1)


def pose_loss(pred_r, pred_t, pred_c, target, model_points, obj, points, w, refine, sym_list):
    """One frame.  pred_r (N,4), pred_t (N,3), pred_c (N,1)|(N,), target/model_points (M,3),
    points (N,3).  Returns loss, dis_sel, new_points (N,3), new_target (M,3), extras dict."""
    n = pred_r.shape[0]
    m = model_points.shape[0]
    c = pred_c.reshape(n)
    q = pred_r / pred_r.norm(dim=1, keepdim=True)
    rot = quat_to_rot(q)                                            # (N, 3, 3)
    pred = torch.einsum("nck,jk->njc", rot, model_points) + (points + pred_t)[:, None, :]
    tgt = target[None].expand(n, m, 3)
    nn_idx = None
    if (not refine) and int(obj) in sym_list:
        nn_idx = nearest_index(target, pred.reshape(-1, 3))
        tgt = target[nn_idx].reshape(n, m, 3)
    dis = (pred - tgt).norm(dim=2).mean(dim=1)                      # (N,)
    loss = (dis * c - w * torch.log(c)).mean()
    which = int(torch.argmax(c))
    t_sel = (pred_t[which] + points[which]).detach()
    r_sel = rot[which].detach()
    new_points = (points - t_sel) @ r_sel
    new_target = (target - t_sel) @ r_sel
    return loss, dis[which], new_points.detach(), new_target.detach(), {
        "dis": dis, "which_max": which, "nn_idx": nn_idx, "R": r_sel, "t": t_sel}


def refine_loss(pred_r, pred_t, target, model_points, obj, points, sym_list):
    """One frame, one pose.  pred_r (1,4)|(4,), pred_t (1,3)|(3,).  Returns dis (1,), new_points,
    new_target, extras."""
    q = pred_r.reshape(4)
    q = q / q.norm()
    rot = quat_to_rot(q)
    t = pred_t.reshape(3)
    pred = model_points @ rot.t() + t
    tgt = target
    nn_idx = None
    if int(obj) in sym_list:
        nn_idx = nearest_index(target, pred)
        tgt = target[nn_idx]
    dis = (pred - tgt).norm(dim=1).mean().reshape(1)
    r_d, t_d = rot.detach(), t.detach()
    new_points = (points - t_d) @ r_d
    new_target = (target - t_d) @ r_d
    return dis, new_points.detach(), new_target.detach(), {"nn_idx": nn_idx, "R": r_d, "t": t_d}


# ------------------------------------------------------------------------------------------------
# independent cross-check of the ADD / ADD-S *values* (src/estimation/errors.py:53-84 semantics)
def add_metric(model_points, rot_gt, t_gt, rot_pr, t_pr):
    a = model_points @ rot_gt.T + t_gt
    b = model_points @ rot_pr.T + t_pr
    return float(np.linalg.norm(a - b, axis=1).mean())


def adds_metric(model_points, rot_gt, t_gt, rot_pr, t_pr):
    a = model_points @ rot_gt.T + t_gt
    b = model_points @ rot_pr.T + t_pr
    d = np.linalg.norm(b[:, None, :] - a[None, :, :], axis=2)
    return float(d.min(axis=1).mean())
