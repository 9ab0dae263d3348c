#!/usr/bin/env python
"""Config C5 (BASELINE.json): data-parallel training step, 8 frames per GPU, num_pt_mesh_large = 2300,
one NCCL gradient all-reduce per optimiser step (plr2_b200/train.py; the same record bench.py prints as `train_c5`).

    python tools/train_step.py                                   # 1 GPU
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/train_step.py
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from plr2_b200 import ops  # noqa: E402
from plr2_b200.train import C5TrainStep  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=8)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--knn", default="fast", choices=["exact", "fast"])
    ap.add_argument("--fwd-tc", type=int, default=1,
                    help="1: training forward on the tensor cores too, with the sign-safe ReLU pass (PLR2_PREC_BF16X2_ALL); "
                         "0: FP32 forward, tensor-core backward (PLR2_PREC_BF16X2)")
    ap.add_argument("--overlap", type=int, default=0,
                    help="1: gradient all-reduce inside the backward op, head weights overlapped with the trunk backward")
    ap.add_argument("--graph", type=int, default=1, help="1: replay the whole step as one CUDA graph")
    ap.add_argument("--deal", default="cost", choices=["cost", "contiguous"], help="assignment of the global batch to ranks")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl")
    ts = C5TrainStep(torch.device("cuda", local), world, rank, args.frames,
                     precision=ops.PREC_BF16X2_ALL if args.fwd_tc else ops.PREC_BF16X2,
                     knn_mode=ops.KNN_FAST if args.knn == "fast" else ops.KNN_EXACT, overlap=bool(args.overlap),
                     cuda_graph=bool(args.graph) and not args.overlap, deal=args.deal)
    rec = ts.run(args.steps, args.warmup)
    if rank == 0:
        print(json.dumps(rec), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
