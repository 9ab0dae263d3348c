"""Per-parameter gradient accuracy of the three precision modes against the CPU oracle (config C1 at full size):
fraction of elements within rtol 1e-3 + 1e-4 of the tensor's largest entry (the tests' rule, tests/util.py), the same
fraction at PURE rtol 1e-3 (no floor: elements that sit near zero by cancellation fail it in any FP32 implementation,
the fp32 column shows by how much), and the largest deviation relative to the largest entry.
GPU only.  Usage: python tools/gradcheck_tc.py"""
import sys, numpy as np, torch
sys.path.insert(0, '.')
from oracle import densefusion_oracle as DO
from plr2_b200 import ops, synth
from plr2_b200.lib.loss import Loss
from tests.test_network_gpu import build_posenet
from tests.util import frac_close, tt
shape, obj = synth.LINEMOD, 3
fr = synth.make_frame(shape, 32, obj=obj)
p = {k: tt(v, grad=True) for k, v in synth.make_params("posenet", shape.num_obj, 31).items()}
oimg = tt(fr["out_img"], grad=True)
opr, opt_, opc, oemb = DO.posenet_forward(p, oimg, tt(fr["points"]), tt(fr["choose"]), obj, shape.num_obj)
ol, od, _, _, _ = DO.pose_loss(opr, opt_, opc, tt(fr["target"]), tt(fr["model_points"]), obj, tt(fr["points"]), 0.015, False, shape.sym_list)
ol.backward()
res = {}
for prec in (ops.PREC_FP32, ops.PREC_BF16X2, ops.PREC_BF16X2_ALL):
    net = build_posenet(shape, 31, prec)
    img = tt(fr["out_img"], "cuda", grad=True)
    points, choose = tt(fr["points"], "cuda")[None], tt(fr["choose"], "cuda")[None, None]
    idx = tt(fr["idx"], "cuda").reshape(1, 1)
    pr, pt, pc, emb = net(img[None], points, choose, idx)
    crit = Loss(shape.num_mesh, list(shape.sym_list))
    loss, dis, _, _ = crit(pr, pt, pc, tt(fr["target"], "cuda")[None], tt(fr["model_points"], "cuda")[None], idx, points, 0.015, False)
    loss.backward()
    for k, v in net.named_parameters():
        g = v.grad.detach().cpu().numpy(); o = p[k].grad.numpy()
        atol = 1e-4 * float(np.abs(o).max())
        res.setdefault(k, []).append((frac_close(g, o, 1e-3, atol), frac_close(g, o, 1e-3, 0.0),
                                      float(np.abs(g - o).max() / (np.abs(o).max() + 1e-30))))
    g = img.grad.detach().cpu().numpy(); o = oimg.grad.numpy()
    res.setdefault("g_out_img", []).append((frac_close(g, o, 1e-3, 1e-4 * float(np.abs(o).max())), frac_close(g, o, 1e-3, 0.0),
                                            float(np.abs(g - o).max() / (np.abs(o).max() + 1e-30))))
for k, v in res.items():
    flag = "  <--" if min(x[0] for x in v) < 0.95 else ""
    print("%-18s fp32 %.3f (pure %.3f) %.1e | bwd-tc %.3f (pure %.3f) %.1e | all-tc %.3f (pure %.3f) %.1e%s"
          % ((k,) + v[0] + v[1] + v[2] + (flag,)))
worst = {m: min(v[i][0] for v in res.values()) for i, m in enumerate(("fp32", "bwd-tc", "all-tc"))}
print("smallest with-floor fraction over all tensors:", {k: round(x, 4) for k, x in worst.items()})
