#!/usr/bin/env python
"""Config C5 (BASELINE.json): data-parallel training step, 8 frames per GPU, num_pt_mesh_large = 2300,
one NCCL gradient all-reduce per optimiser step.

    python tools/train_step.py                                   # 1 GPU
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/train_step.py

Per step and rank: PoseNet (backbone output given) forward on 8 frames -> Loss (ADD-S for symmetric
classes, N*M^2 = 5.3e9 pairs per such frame) -> backward -> GradBucket.all_reduce (average, Lightning's
accumulate_grad_batches semantics) -> Adam.  Prints one JSON line on rank 0: ms/step (max over ranks,
CUDA events), frames/s over all ranks, the all-reduce share, and a replica-consistency check."""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from plr2_b200 import ops, parallel, synth  # noqa: E402
from plr2_b200.lib.loss import Loss  # noqa: E402
from plr2_b200.lib.network import PoseNet  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=8)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--knn", default="fast", choices=["exact", "fast"])
    ap.add_argument("--fwd-tc", type=int, default=0,
                    help="1: training forward on the tensor cores too (PLR2_PREC_BF16X2_ALL; ReLU masks may differ from FP32 at ~1e-5 of the activations)")
    ap.add_argument("--overlap", type=int, default=0,
                    help="1: gradient all-reduce inside the backward op, head weights overlapped with the trunk backward")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl")
    shp = synth.YCB_REFINE
    net = PoseNet(shp.num_points, shp.num_obj, precision=ops.PREC_BF16X2_ALL if args.fwd_tc else ops.PREC_BF16X2)
    net.cnn = torch.nn.Identity()
    net.load_state_dict({k: torch.from_numpy(v) for k, v in synth.make_params("posenet", shp.num_obj, 1).items()}, strict=False)
    net = net.to(dev)
    crit = Loss(shp.num_mesh, list(shp.sym_list), ops.KNN_FAST if args.knn == "fast" else ops.KNN_EXACT)
    params = net.hot_path_parameters()
    bucket = parallel.GradBucket(params)
    in_op = parallel.InOpGradSync(average=True).attach(net) if (args.overlap and world > 1) else None
    opt = torch.optim.Adam(params, lr=1e-4, fused=True)     # one multi-tensor kernel per step
    frs = [synth.make_frame(shp, 5000 + 131 * rank + i) for i in range(args.frames)]
    st = lambda k: torch.from_numpy(np.stack([f[k] for f in frs])).to(dev)
    out_img, points, choose, obj = st("out_img"), st("points"), st("choose")[:, None], st("idx")[:, None]
    target, model = st("target"), st("model_points")
    ar_ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(args.steps)]

    def step(timed, it=0):
        bucket.zero()
        pr, pt, pc, _ = net(out_img, points, choose, obj)
        loss, dis, _, _ = crit(pr, pt, pc, target, model, obj, points, 0.015, False)
        loss.mean().backward()                       # average over the rank's frames
        if timed:
            ar_ev[it][0].record()
        if in_op is None:
            bucket.all_reduce(average=True)          # then over the ranks (in_op: already done inside the backward)
        if timed:
            ar_ev[it][1].record()
        opt.step()
        return loss

    for _ in range(args.warmup):
        step(False)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for it in range(args.steps):
        loss = step(True, it)                        # no host synchronisation inside the timed region
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    ar_ms = sum(a.elapsed_time(b) for a, b in ar_ev)
    chk = torch.stack([p.detach().double().sum() for p in params]).sum().reshape(1)
    if world > 1:
        t = torch.tensor([ms, ar_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ar_ms = float(t[0]), float(t[1])
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        consistent = bool((hi - lo).abs().item() <= 1e-9 * abs(hi.item()))
    else:
        consistent = True
    if rank == 0:
        print(json.dumps({
            "config": "C5: PoseNet fwd+bwd + Loss (ADD-S, M=2300), %d frames/GPU, grad all-reduce + Adam" % args.frames,
            "n_gpus": world, "steps": args.steps, "ms_per_step": ms / args.steps,
            "frames_per_s": world * args.frames * args.steps / (ms / 1e3),
            "train_forward": "tensor cores (PLR2_PREC_BF16X2_ALL)" if args.fwd_tc else "fp32 kernel, backward on tensor cores (PLR2_PREC_BF16X2)",
            "allreduce_ms_per_step": ar_ms / args.steps, "grad_bucket_mb": bucket.numel * 4 / 1e6,
            "allreduce": "inside the backward op, head weights overlapped with the trunk backward" if in_op else "one flat bucket after the backward",
            "replicas_consistent": consistent, "loss_finite": bool(torch.isfinite(loss).all()),
            "sym_frames_rank0": int(sum(int(f["idx"]) in shp.sym_list for f in frs)),
        }), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
