#!/usr/bin/env python
"""Timing of the loss kernels of the inference chain by batch size and class mix (CUDA events, warm):
Loss(refine=True) = add_dis_kernel + loss_reduce + pose_select, and Loss_refine = refine_loss_kernel.

    python tools/loss_bench.py [--frames 64 512]
"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from plr2_b200 import ops, synth  # noqa: E402
from plr2_b200.lib.loss import Loss  # noqa: E402
from plr2_b200.lib.loss_refiner import Loss_refine  # noqa: E402


def ev(fn, reps=20):
    """GPU time per call in us: the call is captured in a CUDA graph and replayed (no host launch overhead in the number)."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        g.replay()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, nargs="+", default=[64, 512])
    args = ap.parse_args()
    shp = synth.YCB_REFINE
    dev = "cuda"
    for F in args.frames:
        for mix, objs in (("no sym", [3]), ("all sym", [12]), ("5/16 sym", [12, 3, 0, 19, 1, 2, 4, 15, 5, 6, 7, 18, 8, 9, 10, 20])):
            base = [synth.make_frame(shp, 100 + i, obj=o, with_image=False) for i, o in enumerate(objs)]
            pick = [i % len(base) for i in range(F)]
            st = lambda k: torch.from_numpy(np.stack([base[i][k] for i in pick])).to(dev)
            tgt, mdl, pts, ob = st("target"), st("model_points"), st("points"), st("idx")[:, None]
            rng = np.random.default_rng(0)
            pr = torch.from_numpy(np.stack([base[i]["gt_q"] + 0.05 * rng.standard_normal(4) for i in pick]).astype(np.float32)).to(dev)
            pt = torch.from_numpy(np.stack([base[i]["gt_t"] + 0.005 * rng.standard_normal(3) for i in pick]).astype(np.float32)).to(dev)
            preds = [synth.make_predictions(base[i], 7 + i) for i in range(len(base))]
            P = [torch.from_numpy(np.stack([preds[i][k] for i in pick])).to(dev) for k in range(3)]
            for mode, mname in ((ops.KNN_FAST, "fast"), (ops.KNN_EXACT, "exact")):
                cr = Loss_refine(shp.num_mesh, list(shp.sym_list), mode)
                with torch.no_grad():
                    t = ev(lambda: cr(pr, pt, tgt, mdl, ob, pts))
                nsym = sum(int(base[i]["idx"]) in shp.sym_list for i in pick)
                print(f"F={F:4d} {mix:9s} Loss_refine[{mname}] {t:8.1f} us   ({nsym} sym frames, {nsym * 2300 * 2300 / t / 1e6:.2f} Tpairs/s)", flush=True)
            cl = Loss(shp.num_mesh, list(shp.sym_list))
            with torch.no_grad():
                t = ev(lambda: cl(P[0], P[1], P[2], tgt, mdl, ob, pts, 0.015, True))
            print(f"F={F:4d} {mix:9s} Loss(refine=True)   {t:8.1f} us   ({F * 1000 * 2300 / t / 1e6:.2f} T hypothesis-points/s)", flush=True)


if __name__ == "__main__":
    main()
