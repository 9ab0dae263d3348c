"""GPU: kernels (1)+(2) — PoseNet / PoseRefineNet drop-ins (forward and backward) against the
reference's golden outputs and the oracle, plus the GEMM entry point against numpy."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import densefusion_oracle as DO
from plr2_b200 import _cabi, ops, synth
from plr2_b200.lib.loss import Loss
from plr2_b200.lib.loss_refiner import Loss_refine
from plr2_b200.lib.network import PoseNet, PoseRefineNet
from tests.util import assert_close, load_golden, shape_of, tt

pytestmark = pytest.mark.gpu


def build_posenet(shp, seed, precision):
    net = PoseNet(shp.num_points, shp.num_obj, precision=precision)
    net.cnn = torch.nn.Identity()        # the `img` argument then carries the backbone output
    net.load_state_dict({k: tt(v) for k, v in synth.make_params("posenet", shp.num_obj, seed).items()}, strict=False)
    return net.cuda()


def build_refiner(shp, seed, precision):
    net = PoseRefineNet(shp.num_points, shp.num_obj, precision=precision)
    net.load_state_dict({k: tt(v) for k, v in synth.make_params("refiner", shp.num_obj, seed).items()})
    return net.cuda()


# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("M,N,K,layout", [(1, 1, 1, "nt"), (130, 70, 33, "nt"), (257, 129, 64, "nt"),
                                          (1000, 640, 384, "nt"), (300, 96, 200, "nn"), (77, 130, 1000, "tn"),
                                          (128, 256, 4096, "tn_split"), (5, 1920, 1024, "nt")])
def test_gemm_f32_layouts_and_epilogues(M, N, K, layout):
    rng = np.random.default_rng(M * 3 + N * 5 + K)
    A = rng.standard_normal((M, K)).astype(np.float32)
    B = rng.standard_normal((K, N)).astype(np.float32)
    bias = rng.standard_normal(N).astype(np.float32)
    d = _cabi.GemmDesc()
    keep = []

    def dev(a):
        t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
        keep.append(t)
        return t

    if layout == "nt":       # A [M,K] k-contig, W [N,K] k-contig
        a, b = dev(A), dev(B.T)
        d.A, d.sam, d.sak, d.B, d.sbk, d.sbn = a.data_ptr(), K, 1, b.data_ptr(), 1, K
    elif layout == "nn":     # A [M,K], B [K,N] n-contig
        a, b = dev(A), dev(B)
        d.A, d.sam, d.sak, d.B, d.sbk, d.sbn = a.data_ptr(), K, 1, b.data_ptr(), N, 1
    else:                    # A stored [K,M] m-contig, B [K,N]
        a, b = dev(A.T), dev(B)
        d.A, d.sam, d.sak, d.B, d.sbk, d.sbn = a.data_ptr(), 1, M, b.data_ptr(), N, 1
    C0 = rng.standard_normal((M, N)).astype(np.float32)
    c = dev(C0)
    d.C, d.ldc, d.M, d.N, d.K = c.data_ptr(), N, M, N, K
    want = A.astype(np.float64) @ B.astype(np.float64)
    if layout == "tn_split":
        d.accumulate, d.split_k = 1, 7
        want = want + C0
    elif layout == "tn":
        d.accumulate = 1
        want = want + C0
    else:
        d.bias, d.relu = dev(bias).data_ptr(), 1
        want = np.maximum(want + bias, 0)
    rc = _cabi.lib().plr2_gemm_f32(ctypes.byref(d), torch.cuda.current_stream().cuda_stream)
    _cabi.check(rc, "plr2_gemm_f32")
    torch.cuda.synchronize()
    assert_close(c.cpu().numpy(), want, "gemm", rtol=1e-4, atol=1e-4 * np.sqrt(K), frac=1.0)


# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["posenet_linemod_nonsym", "posenet_ycb_sym", "posenet_ycb_refine"])
@pytest.mark.parametrize("precision", [ops.PREC_FP32, ops.PREC_BF16X2, ops.PREC_BF16X2_ALL])
def test_posenet_forward_backward_matches_reference_golden(name, precision):
    g = load_golden(name)
    shp = shape_of(g)
    net = build_posenet(shp, int(g["seed"]), precision)
    fr = synth.make_frame(shp, int(g["seed"]) + 1, obj=int(g["obj"]))
    img = tt(fr["out_img"], "cuda")[None]
    points = tt(fr["points"], "cuda")[None]
    choose = tt(fr["choose"], "cuda")[None, None]
    idx = tt(fr["idx"], "cuda").reshape(1, 1)
    pr, pt, pc, emb = net(img, points, choose, idx)
    assert tuple(pr.shape) == (1, shp.num_points, 4) and tuple(pt.shape) == (1, shp.num_points, 3)
    assert tuple(pc.shape) == (1, shp.num_points, 1) and tuple(emb.shape) == (1, 32, shp.num_points)
    assert not emb.requires_grad
    pr.retain_grad(); pt.retain_grad(); pc.retain_grad()
    crit = Loss(shp.num_mesh, list(shp.sym_list))
    loss, dis, npts, ntgt = crit(pr, pt, pc, tt(fr["target"], "cuda")[None], tt(fr["model_points"], "cuda")[None],
                                 idx, points, float(g["w"]), bool(g["refine"]))
    loss.backward()
    c = lambda t: t.detach().cpu().numpy()
    assert np.array_equal(c(emb), g["emb"])
    assert_close(c(pr), g["pred_r"], "pred_r")
    assert_close(c(pt), g["pred_t"], "pred_t")
    assert_close(c(pc), g["pred_c"], "pred_c")
    assert_close(c(loss), g["loss"], "loss", frac=1.0)
    assert_close(c(dis), g["dis"], "dis", frac=1.0)
    assert_close(c(npts), g["new_points"], "new_points")
    assert_close(c(ntgt), g["new_target"], "new_target")
    for k, p in net.named_parameters():
        assert p.grad is not None, k
        assert_close(p.grad.norm().item(), g["gnorm." + k], "gnorm." + k, rtol=2e-3, frac=1.0)
        assert_close(c(p.grad).reshape(-1)[::97], g["gsub." + k], "gsub." + k)


@pytest.mark.parametrize("name", ["refiner_ycb_sym", "refiner_ycb_nonsym"])
@pytest.mark.parametrize("precision", [ops.PREC_FP32, ops.PREC_BF16X2])
def test_refiner_chain_matches_reference_golden(name, precision):
    g = load_golden(name)
    shp = shape_of(g)
    net = build_refiner(shp, int(g["seed"]), precision)
    fr = synth.make_frame(shp, int(g["seed"]) + 1, obj=int(g["obj"]))
    emb = tt(fr["out_img"].reshape(32, -1)[:, fr["choose"]], "cuda")[None]
    crit = Loss_refine(shp.num_mesh, list(shp.sym_list))
    idx = tt(fr["idx"], "cuda").reshape(1, 1)
    pts, tgt = tt(g["in_points"], "cuda")[None], tt(g["in_target"], "cuda")[None]
    c = lambda t: t.detach().cpu().numpy()
    for ite in range(2):       # tools/train.py:141-145
        pr, pt = net(pts, emb, idx)
        assert tuple(pr.shape) == (1, 4) and tuple(pt.shape) == (1, 3)
        dis, pts, tgt = crit(pr, pt, tgt, tt(fr["model_points"], "cuda")[None], idx, pts)
        dis.backward()
        assert_close(c(pr), g[f"it{ite}.pred_r"], "pred_r", frac=1.0)
        assert_close(c(pt), g[f"it{ite}.pred_t"], "pred_t", frac=1.0)
        assert_close(c(dis), g[f"it{ite}.dis"], "dis", frac=1.0, rtol=2e-3)
        assert_close(c(pts), g[f"it{ite}.new_points"], "new_points", rtol=2e-3)
        assert_close(c(tgt), g[f"it{ite}.new_target"], "new_target", rtol=2e-3)
    for k, p in net.named_parameters():      # accumulated over both iterations
        assert_close(p.grad.norm().item(), g["gnorm." + k], "gnorm." + k, rtol=5e-3, frac=1.0)


@pytest.mark.parametrize("precision", [ops.PREC_FP32, ops.PREC_BF16X2])
@pytest.mark.parametrize("shape,obj", [(synth.LINEMOD, 3), (synth.YCB, 12)])
def test_posenet_full_size_vs_oracle(shape, obj, precision):
    """BASELINE configs C1 / C2 at full size (N = 500 / 1000): forward + loss + all gradients, with the
    contractions on the FP32 SIMT kernel and on the tensor cores (packed split-bf16 operands)."""
    net = build_posenet(shape, 31, precision)
    fr = synth.make_frame(shape, 32, obj=obj)
    img = tt(fr["out_img"], "cuda", grad=True)
    points, choose = tt(fr["points"], "cuda")[None], tt(fr["choose"], "cuda")[None, None]
    idx = tt(fr["idx"], "cuda").reshape(1, 1)
    pr, pt, pc, emb = net(img[None], points, choose, idx)
    crit = Loss(shape.num_mesh, list(shape.sym_list))
    loss, dis, _, _ = crit(pr, pt, pc, tt(fr["target"], "cuda")[None], tt(fr["model_points"], "cuda")[None], idx,
                           points, 0.015, False)
    loss.backward()
    p = {k: tt(v, grad=True) for k, v in synth.make_params("posenet", shape.num_obj, 31).items()}
    oimg = tt(fr["out_img"], grad=True)
    opr, opt_, opc, oemb = DO.posenet_forward(p, oimg, tt(fr["points"]), tt(fr["choose"]), obj, shape.num_obj)
    ol, od, _, _, _ = DO.pose_loss(opr, opt_, opc, tt(fr["target"]), tt(fr["model_points"]), obj, tt(fr["points"]),
                                   0.015, False, shape.sym_list)
    ol.backward()
    c = lambda t: t.detach().cpu().numpy()
    assert_close(c(pr)[0], c(opr), "pred_r")
    assert_close(c(pt)[0], c(opt_), "pred_t")
    assert_close(c(pc)[0], c(opc), "pred_c")
    assert_close(c(loss), ol.item(), "loss", frac=1.0)
    assert_close(c(img.grad), c(oimg.grad), "g_out_img", frac=0.99)
    for k, v in net.named_parameters():
        assert_close(c(v.grad), c(p[k].grad), "grad " + k)


def test_batched_frames_equal_single_frame_calls():
    shp = synth.Shape(200, 21, 100, 24, 24, tuple(synth.YCB_SYM))
    net = build_posenet(shp, 5, ops.PREC_FP32)
    frs = [synth.make_frame(shp, 60 + i) for i in range(4)]
    st = lambda key: tt(np.stack([f[key] for f in frs]), "cuda")
    with torch.no_grad():
        pr, pt, pc, emb = net(st("out_img"), st("points"), st("choose")[:, None], st("idx")[:, None])
        for i in range(4):
            a = net(st("out_img")[i:i + 1], st("points")[i:i + 1], st("choose")[i:i + 1, None], st("idx")[i:i + 1, None])
            for x, y in zip(a, (pr, pt, pc, emb)):
                assert torch.equal(x[0], y[i])


@pytest.mark.parametrize("frames,tc_prec", [(6, ops.PREC_BF16X2), (9, ops.PREC_BF16X2), (9, ops.PREC_BF16X2_ALL)])
def test_batched_training_tensor_core_equals_fp32_path(frames, tc_prec):
    """Six / nine frames of 200 points (1200 / 1800 rows: K tail of the weight gradients read MN-major, split-K,
    r/t/c head batches split once, dY planes shared by dW and dX; nine frames put the K = frames contractions on the
    tensor cores too) through forward + Loss + backward: gradients of the tensor-core training paths against the
    FP32 SIMT path."""
    shp = synth.Shape(200, 21, 100, 24, 24, tuple(synth.YCB_SYM))
    frs = [synth.make_frame(shp, 90 + i, obj=o) for i, o in enumerate([3, 12, 0, 20, 7, 15, 1, 19, 5][:frames])]
    st = lambda key: tt(np.stack([f[key] for f in frs]), "cuda")
    grads = {}
    for prec in (ops.PREC_FP32, tc_prec):
        net = build_posenet(shp, 9, prec)
        img = st("out_img").requires_grad_(True)
        pr, pt, pc, _ = net(img, st("points"), st("choose")[:, None], st("idx")[:, None])
        crit = Loss(shp.num_mesh, list(shp.sym_list))
        loss, _, _, _ = crit(pr, pt, pc, st("target"), st("model_points"), st("idx")[:, None], st("points"), 0.015, False)
        loss.mean().backward()
        grads[prec] = ({k: v.grad.detach().cpu().numpy() for k, v in net.named_parameters() if v.grad is not None},
                       img.grad.detach().cpu().numpy(), pr.detach().cpu().numpy())
    ga, gb = grads[ops.PREC_FP32], grads[tc_prec]
    assert_close(gb[2], ga[2], "pred_r")
    assert_close(gb[1], ga[1], "g_out_img", frac=0.99)
    assert set(ga[0]) == set(gb[0]) and len(ga[0]) >= 30
    # Tensor-core FORWARD (PLR2_PREC_BF16X2_ALL): activations differ from the FP32 forward by ~1e-6 relative, so a
    # handful of ReLUs whose pre-activation lies within that distance of zero take the other branch (the sign-safe
    # pass guarantees the sign for the inputs it was given, not for the FP32 path's inputs).  ONE such unit moves the
    # affected weight-gradient entries by ~1/rows (5.6e-4 at 1800 rows): measured 0.87-1.00 of the elements within
    # rtol 1e-3 and at most 1.4e-3 of the tensor's largest entry; the FP32-forward mode is at 1.0000 / 1e-5.
    frac, worst = (0.8, 4e-3) if tc_prec == ops.PREC_BF16X2_ALL else (0.999, 5e-5)
    for k in ga[0]:
        assert_close(gb[0][k], ga[0][k], "grad " + k, frac=frac)
        dev = float(np.abs(gb[0][k] - ga[0][k]).max() / np.abs(ga[0][k]).max())
        assert dev <= worst, (k, dev)


def test_weight_update_is_picked_up():
    shp = synth.Shape(64, 13, 50, 16, 16, tuple(synth.LINEMOD_SYM))
    net = build_posenet(shp, 7, ops.PREC_FP32)
    fr = synth.make_frame(shp, 8)
    args = (tt(fr["out_img"], "cuda")[None], tt(fr["points"], "cuda")[None], tt(fr["choose"], "cuda")[None, None],
            tt(fr["idx"], "cuda").reshape(1, 1))
    with torch.no_grad():
        a = net(*args)[0].clone()
        net.conv4_r.bias.add_(1.0)
        b = net(*args)[0]
    assert torch.allclose(b, a + 1.0, atol=1e-5)


# ---------------------------------------------------------------------------------------------
# tensor-core (tcgen05 split-bf16) inference path
@pytest.mark.parametrize("name", ["posenet_linemod_nonsym", "posenet_ycb_sym", "posenet_ycb_refine"])
def test_posenet_inference_tensor_core_matches_reference_golden(name):
    g = load_golden(name)
    shp = shape_of(g)
    net = build_posenet(shp, int(g["seed"]), ops.PREC_BF16X2)
    fr = synth.make_frame(shp, int(g["seed"]) + 1, obj=int(g["obj"]))
    points = tt(fr["points"], "cuda")[None]
    idx = tt(fr["idx"], "cuda").reshape(1, 1)
    with torch.no_grad():
        (pr, pt, pc, emb), _, n_tc, _ = ops.tensor_core_profile(
            lambda: net(tt(fr["out_img"], "cuda")[None], points, tt(fr["choose"], "cuda")[None, None], idx))
        assert n_tc == (7 if shp.num_points >= 128 else 0), "PoseNet inference must run its 7 tcgen05 GEMM launches"
        crit = Loss(shp.num_mesh, list(shp.sym_list))
        loss, dis, npts, ntgt = crit(pr, pt, pc, tt(fr["target"], "cuda")[None], tt(fr["model_points"], "cuda")[None],
                                     idx, points, float(g["w"]), bool(g["refine"]))
    c = lambda t: t.detach().cpu().numpy()
    assert np.array_equal(c(emb), g["emb"])
    assert_close(c(pr), g["pred_r"], "pred_r")
    assert_close(c(pt), g["pred_t"], "pred_t")
    assert_close(c(pc), g["pred_c"], "pred_c")
    assert_close(c(loss), g["loss"], "loss", frac=1.0)
    assert_close(c(dis), g["dis"], "dis", frac=1.0)
    assert_close(c(npts), g["new_points"], "new_points")
    assert_close(c(ntgt), g["new_target"], "new_target")


@pytest.mark.parametrize("name", ["refiner_ycb_sym", "refiner_ycb_nonsym"])
def test_refiner_inference_tensor_core_matches_reference_golden(name):
    g = load_golden(name)
    shp = shape_of(g)
    net = build_refiner(shp, int(g["seed"]), ops.PREC_BF16X2)
    fr = synth.make_frame(shp, int(g["seed"]) + 1, obj=int(g["obj"]))
    emb = tt(fr["out_img"].reshape(32, -1)[:, fr["choose"]], "cuda")[None]
    idx = tt(fr["idx"], "cuda").reshape(1, 1)
    with torch.no_grad():
        (pr, pt), _, n_tc, _ = ops.tensor_core_profile(lambda: net(tt(g["in_points"], "cuda")[None], emb, idx))
    assert n_tc == 4, "PoseRefineNet inference must run its 4 tcgen05 GEMM launches"
    assert_close(pr.cpu().numpy(), g["it0.pred_r"], "pred_r", frac=1.0)
    assert_close(pt.cpu().numpy(), g["it0.pred_t"], "pred_t", frac=1.0)


@pytest.mark.parametrize("frames", [1, 5])
def test_tensor_core_full_size_vs_oracle_and_fp32_path(frames):
    """YCB shape (N = 1000, 21 classes): tensor-core inference vs the oracle, per element, and
    identical per frame whether batched or not."""
    shp = synth.YCB
    net = build_posenet(shp, 41, ops.PREC_BF16X2)
    frs = [synth.make_frame(shp, 70 + i) for i in range(frames)]
    st = lambda key: tt(np.stack([f[key] for f in frs]), "cuda")
    with torch.no_grad():
        (pr, pt, pc, emb), _, n_tc, flops = ops.tensor_core_profile(
            lambda: net(st("out_img"), st("points"), st("choose")[:, None], st("idx")[:, None]))
    assert n_tc == 7 and flops > 3.5e9 * frames
    p = {k: tt(v) for k, v in synth.make_params("posenet", shp.num_obj, 41).items()}
    for i, fr in enumerate(frs):
        with torch.no_grad():
            opr, opt_, opc, oemb = DO.posenet_forward(p, tt(fr["out_img"]), tt(fr["points"]), tt(fr["choose"]),
                                                      int(fr["idx"]), shp.num_obj)
        assert_close(pr[i].cpu().numpy(), opr.numpy(), "pred_r")
        assert_close(pt[i].cpu().numpy(), opt_.numpy(), "pred_t")
        assert_close(pc[i].cpu().numpy(), opc.numpy(), "pred_c")
        assert np.array_equal(emb[i].cpu().numpy(), oemb.numpy())
    if frames > 1:
        with torch.no_grad():
            a = net(st("out_img")[2:3], st("points")[2:3], st("choose")[2:3, None], st("idx")[2:3, None])
        # the fused average pool sums in 32-row groups, so the summation order (not the math) depends on
        # where the frame starts in the batch: equal to fp32 rounding, not bit for bit
        for x, y in zip(a[:3], (pr, pt, pc)):
            assert_close(x[0].cpu().numpy(), y[2].cpu().numpy(), "batched vs single", rtol=1e-4, frac=1.0)


def test_pipeline_refined_pose_matches_oracle_composition():
    """Kernel (5): the refine chain stays on the device and the per-stage poses are composed there
    (R <- R Rd, t <- R td + t).  Against the oracle chain of tools/train.py:182-188 for 3 frames."""
    from plr2_b200.pipeline import DenseFusionPipeline, INPUT_KEYS
    shp = synth.Shape(1000, 21, 400, 40, 40, tuple(synth.YCB_SYM))
    pipe = DenseFusionPipeline(shp.num_points, shp.num_obj, shp.num_mesh, shp.sym_list, device="cuda:0",
                               precision=ops.PREC_BF16X2, chunk_frames=2, refine_iters=2)
    pn = synth.make_params("posenet", shp.num_obj, 31)
    rf = synth.make_params("refiner", shp.num_obj, 32)
    pipe.load_params(pn, rf)
    frs = [synth.make_frame(shp, 300 + i, obj=o) for i, o in enumerate([12, 3, 19])]
    st = lambda k: tt(np.stack([f[k] for f in frs]), "cuda")
    work = {"out_img": st("out_img"), "points": st("points"), "choose": st("choose")[:, None], "obj": st("idx")[:, None],
            "target": st("target"), "model_points": st("model_points")}
    dis, pose = pipe.infer(*(work[k] for k in INPUT_KEYS), return_pose=True)
    p = {k: tt(v) for k, v in pn.items()}
    r = {k: tt(v) for k, v in rf.items()}
    for i, fr in enumerate(frs):
        obj = int(fr["idx"])
        with torch.no_grad():
            pr, pt, pc, emb = DO.posenet_forward(p, tt(fr["out_img"]), tt(fr["points"]), tt(fr["choose"]), obj, shp.num_obj)
            _, d, npts, ntgt, ex = DO.pose_loss(pr, pt, pc, tt(fr["target"]), tt(fr["model_points"]), obj, tt(fr["points"]),
                                                0.015, True, shp.sym_list)
            R, t = ex["R"].double(), ex["t"].double()
            for _ in range(2):
                rr, rt = DO.refiner_forward(r, npts, emb, obj, shp.num_obj)
                d, npts, ntgt, ex = DO.refine_loss(rr, rt, ntgt, tt(fr["model_points"]), obj, npts, shp.sym_list)
                t = R @ ex["t"].double() + t
                R = R @ ex["R"].double()
        assert_close(dis[i].cpu().numpy(), d.numpy(), "final dis", rtol=2e-3, frac=1.0)
        assert_close(pose[i, :9].cpu().numpy(), R.numpy().reshape(-1), "composed R", frac=1.0)
        assert_close(pose[i, 9:].cpu().numpy(), t.numpy(), "composed t", frac=1.0)
        # the composed pose maps the model onto the (re-assembled) target cloud about as well as `dis` says
        m = tt(fr["model_points"]).double()
        pred = m @ R.T + t
        if obj not in shp.sym_list:
            assert abs(float((pred - tt(fr["target"]).double()).norm(dim=1).mean()) - float(d)) < 1e-4


def test_refiner_next_iteration_reuses_the_emb_branch_bit_identically():
    """forward_next_iteration (refine iteration >= 2, same emb) must give exactly what a full forward gives, and
    must fall back to the full forward when the call does not continue the previous one."""
    g = load_golden("refiner_ycb_sym")
    shp = shape_of(g)
    net = build_refiner(shp, int(g["seed"]), ops.PREC_BF16X2)
    frames = 3
    gen = torch.Generator().manual_seed(77)
    emb = torch.randn(frames, 32, shp.num_points, generator=gen).cuda()
    emb_other = torch.randn(frames, 32, shp.num_points, generator=gen).cuda()
    x1 = (torch.randn(frames, shp.num_points, 3, generator=gen) * 0.05).cuda()
    x2 = (torch.randn(frames, shp.num_points, 3, generator=gen) * 0.05).cuda()
    idx = torch.tensor([[3], [12], [20]], device="cuda")
    with torch.no_grad():
        full2 = [t.clone() for t in net(x2, emb, idx)]
        full2_other = [t.clone() for t in net(x2, emb_other, idx)]
        net(x1, emb, idx)                                           # iteration 1
        (r, t), _, n_tc, _ = ops.tensor_core_profile(lambda: net.forward_next_iteration(x2, emb, idx))
        assert n_tc == 4
        assert torch.equal(r, full2[0]) and torch.equal(t, full2[1])
        # different emb tensor: not a continuation -> full forward, still right
        r, t = net.forward_next_iteration(x2, emb_other, idx)
        assert torch.equal(r, full2_other[0]) and torch.equal(t, full2_other[1])
        # emb modified in place between the calls: version counter differs -> full forward
        net(x1, emb, idx)
        emb.mul_(0.5)
        r, t = net.forward_next_iteration(x2, emb, idx)
        ref = net(x2, emb, idx)
        assert torch.equal(r, ref[0]) and torch.equal(t, ref[1])


# ---------------------------------------------------------------------------------------------
# boundary behaviour: input-range flags, packed-weight cache, backward guards
def _small_posenet_inputs(shp, frames=2):
    frs = [synth.make_frame(shp, 9100 + i, obj=3 + i) for i in range(frames)]
    st = lambda k: tt(np.stack([f[k] for f in frs]), "cuda")
    return st("out_img"), st("points"), st("choose")[:, None], st("idx")[:, None]


@pytest.mark.parametrize("precision", [ops.PREC_FP32, ops.PREC_BF16X2])
def test_out_of_range_choose_and_class_are_reported(precision):
    """torch.gather / index_select raise on a bad index (lib/network.py:102,123-126); the kernels clamp and flag it:
    with check_inputs the drop-in raises IndexError too, without it the call goes through (clamped)."""
    shp = synth.Shape(256, 21, 50, 20, 20, tuple(synth.YCB_SYM))
    net = build_posenet(shp, 5, precision)
    img, pts, ch, ob = _small_posenet_inputs(shp)
    net.check_inputs = True
    with torch.no_grad():
        net(img, pts, ch, ob)                                      # clean inputs: no error
        bad = ch.clone(); bad[1, 0, 17] = shp.height * shp.width   # one past the end
        with pytest.raises(IndexError, match="choose"):
            net(img, pts, bad, ob)
        bad[1, 0, 17] = -1
        with pytest.raises(IndexError, match="choose"):
            net(img, pts, bad, ob)
        with pytest.raises(IndexError, match="class index"):
            net(img, pts, ch, torch.tensor([[3], [shp.num_obj]], device="cuda"))
        net(img, pts, ch, ob)                                      # flags are per forward: clean again
        net.check_inputs = False
        out = net(img, pts, bad, ob)                               # release behaviour: clamped, finite
        assert all(bool(torch.isfinite(t).all()) for t in out)
    ref = build_refiner(shp, 6, precision)
    ref.check_inputs = True
    emb = torch.randn(2, 32, shp.num_points, device="cuda")
    with torch.no_grad():
        ref(pts, emb, ob)
        with pytest.raises(IndexError, match="class index"):
            ref(pts, emb, torch.tensor([[-2], [3]], device="cuda"))


def test_packed_weight_cache_and_backward_guards():
    shp = synth.Shape(256, 21, 50, 20, 20, tuple(synth.YCB_SYM))
    net = build_posenet(shp, 7, ops.PREC_BF16X2)
    img, pts, ch, ob = _small_posenet_inputs(shp)
    with torch.no_grad():
        a = net(img, pts, ch, ob)[0].clone()
        net.conv4_r.bias.data.add_(0.5)              # .data: no version bump -> stale packed weights until told
        stale = net(img, pts, ch, ob)[0].clone()
        net.invalidate_packed_weights()
        fresh = net(img, pts, ch, ob)[0].clone()
        assert torch.equal(a, stale)
        assert_close((fresh - a).cpu().numpy(), np.full(tuple(a.shape), 0.5, np.float32), "bias shift", rtol=1e-4, frac=1.0)
        net.conv4_r.bias.add_(0.25)                  # in-place on the parameter: detected by its version counter
        assert_close((net(img, pts, ch, ob)[0] - a).cpu().numpy(), np.full(tuple(a.shape), 0.75, np.float32), "bias shift",
                     rtol=1e-4, frac=1.0)
    pr, pt, pc, _ = net(img, pts, ch, ob)
    loss = pr.sum() + pt.sum() + pc.sum()
    loss.backward(retain_graph=True)
    with pytest.raises(RuntimeError, match="second time"):
        loss.backward()
    pr, pt, pc, _ = net(img, pts, ch, ob)
    with torch.no_grad():
        net.conv1_r.weight.mul_(1.01)                # modified between forward and backward
    with pytest.raises(RuntimeError, match="modified by an inplace operation"):
        (pr.sum() + pt.sum() + pc.sum()).backward()


def test_in_place_gradient_accumulation_equals_autograd_accumulation():
    """accumulate_grads_in_place: ONE kernel adds the packed gradient onto the existing .grad tensors; same bits as
    autograd's per-parameter AccumulateGrad, over two backward passes (tools/train.py:150-155 accumulates frames)."""
    shp = synth.Shape(256, 21, 50, 20, 20, tuple(synth.YCB_SYM))
    img, pts, ch, ob = _small_posenet_inputs(shp)
    grads = []
    for in_place in (False, True):
        net = build_posenet(shp, 7, ops.PREC_BF16X2)
        net.accumulate_grads_in_place(in_place)
        params = net.hot_path_parameters()
        for p in params:
            p.grad = torch.full_like(p, 0.125)       # something already accumulated
        for rep in range(2):
            pr, pt, pc, _ = net(img, pts, ch, ob)
            (pr.sum() * (1 + rep) + pt.sum() - pc.sum()).backward()
        grads.append([p.grad.clone() for p in params])
    for a, b in zip(*grads):
        assert torch.equal(a, b)
    # parameters without a .grad fall back to the ordinary path
    net = build_posenet(shp, 7, ops.PREC_BF16X2).accumulate_grads_in_place(True)
    pr, pt, pc, _ = net(img, pts, ch, ob)
    (pr.sum() + pt.sum() - pc.sum()).backward()
    assert all(p.grad is not None for p in net.hot_path_parameters())


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_one_process_two_devices():
    """Function attributes (232 KB of dynamic shared memory for the tcgen05 GEMM, the k-NN tile) and occupancy figures
    are per DEVICE: the same process must be able to run the whole pipeline on cuda:0 and then on cuda:1."""
    from plr2_b200.lib.knn import knn_indices
    from plr2_b200.pipeline import DenseFusionPipeline, INPUT_KEYS
    shp = synth.Shape(1000, 21, 2300, 40, 40, tuple(synth.YCB_SYM))
    pn = synth.make_params("posenet", shp.num_obj, 31)
    rf = synth.make_params("refiner", shp.num_obj, 32)
    frs = [synth.make_frame(shp, 300 + i, obj=o) for i, o in enumerate([12, 3, 19])]
    res = []
    for dev in ("cuda:0", "cuda:1"):
        pipe = DenseFusionPipeline(shp.num_points, shp.num_obj, shp.num_mesh, shp.sym_list, device=dev, chunk_frames=4)
        pipe.load_params(pn, rf)
        st = lambda k: tt(np.stack([f[k] for f in frs]), dev)
        work = {"out_img": st("out_img"), "points": st("points"), "choose": st("choose")[:, None], "obj": st("idx")[:, None],
                "target": st("target"), "model_points": st("model_points")}
        res.append(pipe.infer(*(work[k] for k in INPUT_KEYS)).cpu())
        rng = np.random.default_rng(1)
        ref = torch.from_numpy(rng.standard_normal((1, 3, 5000)).astype(np.float32)).to(dev)     # > 3072 refs: large smem tile
        qry = torch.from_numpy(rng.standard_normal((1, 3, 4000)).astype(np.float32)).to(dev)
        res.append(knn_indices(ref, qry, 1).cpu())
    assert torch.equal(res[0], res[2]) and torch.equal(res[1], res[3])
