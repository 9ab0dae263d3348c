#!/usr/bin/env python
"""Operand-format sweep behind the precision scheme of the tcgen05 GEMM chain.

For every contraction of PoseNet's per-point path (lib/network.py:53-68,107-121) the operands are rounded the way a
given tensor-core format would see them, everything else stays at the baseline format, and the per-point outputs
pred_r / pred_t / pred_c are compared with a float64 evaluation of the same network: fraction of elements within
rtol 1e-3 (pure, and with the atol floor of tests/util.py).  The products of two <= 11-bit-mantissa operands are exact in
FP32, so rounding the operands and multiplying in FP32 (TF32 off) reproduces the tensor-core arithmetic up to the
summation order.  This is evidence tooling (torch ops, library matmul) - not the product path.

Formats (cost in bf16-rate MMAs per algorithmic MAC):
    bf16x1   X bf16,        W bf16            1      kind::f16
    fp16x1   X fp16,        W fp16            1      kind::f16 (11-bit mantissa, same rate as bf16)
    tf32x1   X tf32,        W tf32            2      kind::tf32 runs at half the bf16 rate
    bf16x2   X bf16 hi,     W bf16 hi+lo      2      (the "pooled-only" layers of round 1)
    fp16x2   X fp16 hi,     W fp16 hi+lo      2
    fp16x2a  X fp16 hi+lo,  W fp16 hi         2
    bf16x3   X bf16 hi+lo,  W bf16 hi+lo      3      (round 1 default; lo*lo dropped)
    fp16x3   X fp16 hi+lo,  W fp16 hi+lo      3

    python tools/precision_sweep.py [--frames 8] [--out profiles/r2_precision_sweep.txt]
"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from plr2_b200 import synth  # noqa: E402

COST = {"f32": 0, "bf16x1": 1, "fp16x1": 1, "tf32x1": 2, "bf16x2": 2, "fp16x2": 2, "fp16x2a": 2, "bf16x3": 3, "fp16x3": 3}
# layer group -> MAC per point (PoseNet, hoisted form)
MACS = {"c2": 2 * 64 * 128, "c5": 256 * 512, "c6": 512 * 1024, "h1": 384 * 1920, "h2": 3 * 640 * 256, "h3": 3 * 256 * 128}
GROUPS = ["c2", "c5", "c6", "hb", "h1", "h2", "h3"]
ROUND1 = {"c2": "bf16x3", "c5": "bf16x2", "c6": "bf16x2", "hb": "bf16x3", "h1": "bf16x3", "h2": "bf16x3", "h3": "bf16x3"}


def _rt(x, dt):
    return x.to(dt).to(torch.float32)


def _tf32(x):
    i = x.contiguous().view(torch.int32)
    i = (i + 0x1000) & ~0x1FFF          # cvt.rna.tf32.f32: round to nearest, ties away, 10 mantissa bits
    return i.view(torch.float32)


def quant(x, fmt, is_weight):
    """The value the tensor core effectively multiplies for operand x under `fmt`."""
    if fmt == "f32":
        return x
    if fmt == "tf32x1":
        return _tf32(x)
    dt = torch.bfloat16 if fmt.startswith("bf16") else torch.float16
    hi = _rt(x, dt)
    both = hi + _rt(x - hi, dt)
    if fmt.endswith("x1"):
        return hi
    if fmt.endswith("x2"):
        return both if is_weight else hi
    if fmt.endswith("x2a"):
        return hi if is_weight else both
    return both


def lin(x, w, b, fmt, relu=True):
    """x (R, K) @ w (N, K)^T + b; FP32 accumulate."""
    y = quant(x, fmt, False) @ quant(w, fmt, True).t()
    if b is not None:
        y = y + b
    return torch.relu(y) if relu else y


def posenet(p, out_img, x, choose, obj, num_obj, fmt, dtype=torch.float32):
    """One frame, point-major restatement of lib/network.py:95-132 in the hoisted form the kernels use
    (csrc/posenet.cu): returns pred_r (N,4), pred_t (N,3), pred_c (N,1)."""
    c = lambda t: t.to(dtype)
    W = lambda n: c(p[n + ".weight"].reshape(p[n + ".weight"].shape[0], -1))
    B = lambda n: c(p[n + ".bias"])
    if dtype == torch.float64:
        fmt = {g: "f32" for g in GROUPS}
    emb = c(out_img).reshape(32, -1)[:, choose].t()                       # (N,32)
    x1 = torch.relu(c(x) @ W("feat.conv1").t() + B("feat.conv1"))         # first layer: FP32 SIMT in the product
    e1 = torch.relu(emb @ W("feat.e_conv1").t() + B("feat.e_conv1"))
    x2 = lin(x1, W("feat.conv2"), B("feat.conv2"), fmt["c2"])
    e2 = lin(e1, W("feat.e_conv2"), B("feat.e_conv2"), fmt["c2"])
    pf = torch.cat([x1, e1, x2, e2], 1)                                   # (N,384)
    x5 = lin(pf[:, 128:], W("feat.conv5"), B("feat.conv5"), fmt["c5"])
    x6 = lin(x5, W("feat.conv6"), B("feat.conv6"), fmt["c6"])
    ap = x6.mean(0, keepdim=True)                                         # (1,1024)
    outs = []
    for h, width in (("r", 4), ("t", 3), ("c", 1)):
        w1 = W(f"conv1_{h}")
        hb = lin(ap, w1[:, 384:], B(f"conv1_{h}"), fmt["hb"], relu=False)   # per-frame bias (M = frames GEMM)
        y = torch.relu(quant(pf, fmt["h1"], False) @ quant(w1[:, :384], fmt["h1"], True).t() + hb)
        y = lin(y, W(f"conv2_{h}"), B(f"conv2_{h}"), fmt["h2"])
        y = lin(y, W(f"conv3_{h}"), B(f"conv3_{h}"), fmt["h3"])
        y = y @ W(f"conv4_{h}").t() + B(f"conv4_{h}")                      # FP32 in the epilogue (accumulator registers)
        if h == "c":
            y = torch.sigmoid(y)
        outs.append(y.reshape(-1, num_obj, width)[:, int(obj)])
    return outs


def frac_within(a, b, rtol=1e-3, floor=0.0):
    a, b = a.double().reshape(-1), b.double().reshape(-1)
    atol = floor * float(b.abs().max())
    return float(((a - b).abs() <= atol + rtol * b.abs()).double().mean())


def evaluate(fmt, frames, params, shape, dev):
    """min over frames of the within-rtol fraction per output, pure rtol and with the tests' atol floor (1e-4 max|b|)."""
    res = {k: [] for k in ("r", "t", "c", "r_floor", "t_floor", "c_floor")}
    for fr in frames:
        args = (torch.from_numpy(fr["out_img"]).to(dev), torch.from_numpy(fr["points"]).to(dev),
                torch.from_numpy(fr["choose"]).to(dev), int(fr["idx"]), shape.num_obj)
        with torch.no_grad():
            got = posenet(params, *args, fmt)
            ref = posenet(params, *args, fmt, dtype=torch.float64)
        for k, g, r in zip("rtc", got, ref):
            res[k].append(frac_within(g, r))
            res[k + "_floor"].append(frac_within(g, r, floor=1e-4))
    return {k: (min(v), float(np.mean(v))) for k, v in res.items()}


def mma_cost(fmt):
    return sum(MACS[g] * COST[fmt[g]] for g in MACS) / sum(MACS.values())


def sweep(n_frames=8, dev="cuda", shape=synth.YCB, param_seeds=(1, 7), log=print):
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    rows = []
    for ps in param_seeds:
        params = {k: torch.from_numpy(v).to(dev) for k, v in synth.make_params("posenet", shape.num_obj, ps).items()}
        frames = [synth.make_frame(shape, 3000 + 17 * i, obj=[12, 3, 19, 0, 7, 20, 5, 16][i % 8]) for i in range(n_frames)]

        def run(name, fmt):
            e = evaluate(fmt, frames, params, shape, dev)
            rows.append((ps, name, dict(fmt), mma_cost(fmt), e))
            log("params %d  %-34s MMAs/MAC %.2f  min frac within rtol 1e-3:  r %.4f  t %.4f  c %.4f   (with atol floor: r %.4f t %.4f c %.4f)"
                % (ps, name, mma_cost(fmt), e["r"][0], e["t"][0], e["c"][0], e["r_floor"][0], e["t_floor"][0], e["c_floor"][0]))

        run("fp32 everywhere (summation order only)", {g: "f32" for g in GROUPS})
        run("round-1 scheme", ROUND1)
        base = {g: "bf16x3" for g in GROUPS}
        run("all bf16x3", base)
        for g in GROUPS:
            for f in ("bf16x1", "fp16x1", "tf32x1", "bf16x2", "fp16x2"):
                run(f"{g}={f}, rest bf16x3", dict(base, **{g: f}))
        for f in ("bf16x1", "fp16x1", "tf32x1", "bf16x2", "fp16x2", "fp16x3"):
            run(f"all {f}", {g: f for g in GROUPS})
        f3 = {g: "fp16x3" for g in GROUPS}
        for g in ("h1", "h2", "h3"):
            run(f"{g}=fp16x2a, rest fp16x3", dict(f3, **{g: "fp16x2a"}))
        run("fp16x3, trunk c5,c6 = fp16x2", dict(f3, c5="fp16x2", c6="fp16x2"))
        run("fp16x3, trunk c5,c6 = fp16x1", dict(f3, c5="fp16x1", c6="fp16x1"))
        run("fp16x3, c6 = fp16x1", dict(f3, c6="fp16x1"))
        run("fp16x3, c6 = fp16x1, c5 = fp16x2", dict(f3, c6="fp16x1", c5="fp16x2"))
        run("fp16x3, trunk c2,c5,c6 = fp16x1", dict(f3, c2="fp16x1", c5="fp16x1", c6="fp16x1"))
        run("fp16x3, c5,c6,hb = fp16x1", dict(f3, c5="fp16x1", c6="fp16x1", hb="fp16x1"))
        run("fp16x3, h3 = fp16x2", dict(f3, h3="fp16x2"))
        run("fp16x2 everywhere, c5/c6 fp16x1", dict({g: "fp16x2" for g in GROUPS}, c5="fp16x1", c6="fp16x1"))
        run("fp16x2 heads, fp16x1 trunk (c2,c5,c6)", dict({g: "fp16x2" for g in GROUPS}, c2="fp16x1", c5="fp16x1", c6="fp16x1"))
        run("h1,h2 fp16x1, rest fp16x2", dict({g: "fp16x2" for g in GROUPS}, h1="fp16x1", h2="fp16x1"))
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=8)
    ap.add_argument("--out", default=None)
    ap.add_argument("--device", default="cuda" if torch.cuda.is_available() else "cpu")
    args = ap.parse_args()
    lines = []

    def log(s):
        print(s, flush=True)
        lines.append(s)
    log("# operand-format sweep, PoseNet per-point outputs vs float64, C2 shape (N=1000, 21 classes), %d frames x 2 weight seeds" % args.frames)
    log("# min over frames of the fraction of elements within rtol 1e-3 (pure), then with atol = 1e-4 max|ref| (tests/util.py)")
    log("# device: %s (torch %s; FP32 matmul, TF32 off)" % (args.device, torch.__version__))
    sweep(args.frames, dev=args.device, log=log)
    if args.out:
        with open(args.out, "w") as f:
            f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()


# ---------------------------------------------------------------------------------------------
# PoseRefineNet (lib/network.py:151-206): everything behind conv6 sees the per-frame MEAN of 1000 points, so the
# rounding of the activations averages out - the rounding of the WEIGHTS is the same for every point and does not.
def refiner(p, x, emb, obj, num_obj, fmt, dtype=torch.float32):
    c = lambda t: t.to(dtype)
    W = lambda n: c(p[n + ".weight"].reshape(p[n + ".weight"].shape[0], -1))
    B = lambda n: c(p[n + ".bias"])
    if dtype == torch.float64:
        fmt = {g: "f32" for g in GROUPS}
    x1 = torch.relu(c(x) @ W("feat.conv1").t() + B("feat.conv1"))
    e1 = torch.relu(c(emb).t() @ W("feat.e_conv1").t() + B("feat.e_conv1"))
    x2 = lin(x1, W("feat.conv2"), B("feat.conv2"), fmt["c2"])
    e2 = lin(e1, W("feat.e_conv2"), B("feat.e_conv2"), fmt["c2"])
    pf = torch.cat([x1, e1, x2, e2], 1)
    x5 = lin(pf, W("feat.conv5"), B("feat.conv5"), fmt["c5"])
    x6 = lin(x5, W("feat.conv6"), B("feat.conv6"), fmt["c6"])
    ap = x6.mean(0, keepdim=True)
    out = []
    for h, width in (("r", 4), ("t", 3)):
        y = lin(ap, W(f"conv1_{h}"), B(f"conv1_{h}"), fmt["hb"])           # M = frames tensor-core GEMM
        y = torch.relu(y @ W(f"conv2_{h}").t() + B(f"conv2_{h}"))          # FP32 SIMT
        y = y @ W(f"conv3_{h}").t() + B(f"conv3_{h}")
        out.append(y.reshape(num_obj, width)[int(obj)])
    return out


def sweep_refiner(n_frames=16, dev="cpu", shape=synth.YCB_REFINE, param_seeds=(2, 8), log=print):
    torch.backends.cuda.matmul.allow_tf32 = False
    for ps in param_seeds:
        params = {k: torch.from_numpy(v).to(dev) for k, v in synth.make_params("refiner", shape.num_obj, ps).items()}
        frames = [synth.make_frame(shape, 3500 + 17 * i, obj=[12, 3, 19, 0, 7, 20, 5, 16][i % 8]) for i in range(n_frames)]
        g = torch.Generator().manual_seed(ps)

        def run(name, fmt):
            worst = {"r": 0.0, "t": 0.0}
            for fr in frames:
                x = torch.from_numpy(fr["points"] - fr["points"].mean(0)).to(dev)        # re-centred cloud, like new_points
                emb = torch.from_numpy(fr["out_img"].reshape(32, -1)[:, fr["choose"]]).to(dev)
                with torch.no_grad():
                    got = refiner(params, x, emb, int(fr["idx"]), shape.num_obj, fmt)
                    ref = refiner(params, x, emb, int(fr["idx"]), shape.num_obj, fmt, dtype=torch.float64)
                for k, a, b in zip("rt", got, ref):
                    worst[k] = max(worst[k], float(((a.double() - b).abs() / b.abs().max()).max()))
            log("refiner params %d  %-34s max |err| / max|ref| over %d frames:  pred_r %.2e   pred_t %.2e"
                % (ps, name, n_frames, worst["r"], worst["t"]))
        f3 = {g_: "fp16x3" for g_ in GROUPS}
        run("fp32 everywhere", {g_: "f32" for g_ in GROUPS})
        run("round-1 scheme (bf16; c2,c5,c6 = x2)", dict({g_: "bf16x3" for g_ in GROUPS}, c2="bf16x2", c5="bf16x2", c6="bf16x2"))
        run("fp16; c2,c5,c6 = x2", dict(f3, c2="fp16x2", c5="fp16x2", c6="fp16x2"))
        run("fp16; c2,c5 = x2, c6 = x1", dict(f3, c2="fp16x2", c5="fp16x2", c6="fp16x1"))
        run("fp16; c2,c5,c6 = x1", dict(f3, c2="fp16x1", c5="fp16x1", c6="fp16x1"))
        run("bf16; c6 = x1", dict({g_: "bf16x3" for g_ in GROUPS}, c2="bf16x2", c5="bf16x2", c6="bf16x1"))
