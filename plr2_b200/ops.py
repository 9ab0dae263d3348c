"""autograd Functions over the C-ABI (include/plr2_b200.h).  Frame-batched: every tensor carries a
leading frame dimension F; the reference only ever uses F == 1."""
import torch

from . import _cabi

KNN_EXACT = 0
KNN_FAST = 1

_sym_luts = {}


def sym_lut(sym_list, device):
    """uint8 device table lut[class] = 1 for symmetric classes (replaces the host-side
    `idx[0].item() in sym_list`, lib/loss.py:40, so that no host sync is needed)."""
    key = (tuple(int(s) for s in sym_list), str(device))
    t = _sym_luts.get(key)
    if t is None:
        n = (max(key[0]) + 1) if key[0] else 1
        host = torch.zeros(n, dtype=torch.uint8)
        for s in key[0]:
            if s >= 0:
                host[s] = 1
        t = host.to(device)
        _sym_luts[key] = t
    return t


def _f32c(t):
    return t.detach().to(torch.float32).contiguous()


def _need_cuda(t, what):
    if not t.is_cuda:
        raise _cabi.Plr2Error(f"{what}: CUDA tensors required (plr2_b200 has no CPU fallback)")


class PoseLossFn(torch.autograd.Function):
    """Loss (lib/loss.py:12-69) for F frames.  Returns loss (F,), dis_sel (F,), new_points (F,N,3),
    new_target (F,M,3), dis_all (F,N), which_max (F,) int32."""

    @staticmethod
    def forward(ctx, pred_r, pred_t, pred_c, target, model_points, obj, points, w, refine, lut, knn_mode):
        _need_cuda(pred_r, "PoseLossFn")
        F, N = pred_r.shape[0], pred_r.shape[1]
        M = model_points.shape[1]
        dev = pred_r.device
        pr, pt, pc = _f32c(pred_r), _f32c(pred_t), _f32c(pred_c).reshape(F, N)
        tg, mp, ps = _f32c(target), _f32c(model_points), _f32c(points)
        ob = obj.detach().reshape(F).to(torch.int64).contiguous()
        need_jac = any(ctx.needs_input_grad[:3])
        dis = torch.empty((F, N), dtype=torch.float32, device=dev)
        loss = torch.empty((F,), dtype=torch.float32, device=dev)
        jac_r = torch.empty((F, N, 4), dtype=torch.float32, device=dev) if need_jac else None
        jac_t = torch.empty((F, N, 3), dtype=torch.float32, device=dev) if need_jac else None
        which = torch.empty((F,), dtype=torch.int32, device=dev)
        dis_sel = torch.empty((F,), dtype=torch.float32, device=dev)
        new_points = torch.empty((F, N, 3), dtype=torch.float32, device=dev)
        new_target = torch.empty((F, M, 3), dtype=torch.float32, device=dev)
        L = _cabi.lib()
        with torch.cuda.device(dev):
            st = _cabi.stream_ptr(dev)
            rc = L.plr2_pose_loss_fwd(_cabi.ptr(pr), _cabi.ptr(pt), _cabi.ptr(pc), _cabi.ptr(tg), _cabi.ptr(mp),
                                      _cabi.ptr(ps), _cabi.ptr(ob), _cabi.ptr(lut), int(lut.numel()), F, N, M,
                                      float(w), int(bool(refine)), int(knn_mode), _cabi.ptr(dis), _cabi.ptr(loss),
                                      _cabi.ptr(jac_r), _cabi.ptr(jac_t), st)
            _cabi.check(rc, "plr2_pose_loss_fwd")
            rc = L.plr2_pose_select(_cabi.ptr(pr), _cabi.ptr(pt), _cabi.ptr(pc), _cabi.ptr(dis), _cabi.ptr(tg),
                                    _cabi.ptr(ps), F, N, M, _cabi.ptr(which), _cabi.ptr(dis_sel),
                                    _cabi.ptr(new_points), _cabi.ptr(new_target), None, st)
            _cabi.check(rc, "plr2_pose_select")
        if need_jac:
            ctx.save_for_backward(pc, dis, jac_r, jac_t, which)
        ctx.w = float(w)
        ctx.shapes = (pred_r.shape, pred_t.shape, pred_c.shape)
        ctx.mark_non_differentiable(new_points, new_target, dis, which)
        return loss, dis_sel, new_points, new_target, dis, which

    @staticmethod
    def backward(ctx, g_loss, g_dis_sel, *_):
        pc, dis, jac_r, jac_t, which = ctx.saved_tensors
        F, N = pc.shape
        dev = pc.device
        g_r = torch.empty((F, N, 4), dtype=torch.float32, device=dev)
        g_t = torch.empty((F, N, 3), dtype=torch.float32, device=dev)
        g_c = torch.empty((F, N), dtype=torch.float32, device=dev)
        gl = _f32c(g_loss).reshape(F) if g_loss is not None else None
        gd = _f32c(g_dis_sel).reshape(F) if g_dis_sel is not None else None
        with torch.cuda.device(dev):
            rc = _cabi.lib().plr2_pose_loss_bwd(_cabi.ptr(gl), _cabi.ptr(gd), _cabi.ptr(which), _cabi.ptr(pc),
                                                _cabi.ptr(dis), _cabi.ptr(jac_r), _cabi.ptr(jac_t), F, N, ctx.w,
                                                _cabi.ptr(g_r), _cabi.ptr(g_t), _cabi.ptr(g_c),
                                                _cabi.stream_ptr(dev))
        _cabi.check(rc, "plr2_pose_loss_bwd")
        sr, st_, sc = ctx.shapes
        return (g_r.reshape(sr), g_t.reshape(st_), g_c.reshape(sc), None, None, None, None, None, None, None, None)


class RefineLossFn(torch.autograd.Function):
    """Loss_refine (lib/loss_refiner.py:12-62) for F frames: dis (F,), new_points, new_target."""

    @staticmethod
    def forward(ctx, pred_r, pred_t, target, model_points, obj, points, lut, knn_mode):
        _need_cuda(pred_r, "RefineLossFn")
        F = target.shape[0]
        M, N = model_points.shape[1], points.shape[1]
        dev = pred_r.device
        pr, pt = _f32c(pred_r).reshape(F, 4), _f32c(pred_t).reshape(F, 3)
        tg, mp, ps = _f32c(target), _f32c(model_points), _f32c(points)
        ob = obj.detach().reshape(F).to(torch.int64).contiguous()
        need_jac = any(ctx.needs_input_grad[:2])
        dis = torch.empty((F,), dtype=torch.float32, device=dev)
        jac_r = torch.empty((F, 4), dtype=torch.float32, device=dev) if need_jac else None
        jac_t = torch.empty((F, 3), dtype=torch.float32, device=dev) if need_jac else None
        new_points = torch.empty((F, N, 3), dtype=torch.float32, device=dev)
        new_target = torch.empty((F, M, 3), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = _cabi.lib().plr2_refine_loss_fwd(_cabi.ptr(pr), _cabi.ptr(pt), _cabi.ptr(tg), _cabi.ptr(mp),
                                                  _cabi.ptr(ps), _cabi.ptr(ob), _cabi.ptr(lut), int(lut.numel()),
                                                  F, N, M, int(knn_mode), _cabi.ptr(dis), _cabi.ptr(new_points),
                                                  _cabi.ptr(new_target), _cabi.ptr(jac_r), _cabi.ptr(jac_t), None,
                                                  _cabi.stream_ptr(dev))
        _cabi.check(rc, "plr2_refine_loss_fwd")
        if need_jac:
            ctx.save_for_backward(jac_r, jac_t)
        ctx.shapes = (pred_r.shape, pred_t.shape)
        ctx.mark_non_differentiable(new_points, new_target)
        return dis, new_points, new_target

    @staticmethod
    def backward(ctx, g_dis, *_):
        jac_r, jac_t = ctx.saved_tensors
        g = g_dis.reshape(-1, 1).to(torch.float32)
        return ((g * jac_r).reshape(ctx.shapes[0]), (g * jac_t).reshape(ctx.shapes[1]), None, None, None, None, None, None)
