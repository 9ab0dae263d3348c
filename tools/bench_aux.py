#!/usr/bin/env python
"""Measurement of the widened rows: input-step and metric kernels on one B200 next to the
oracle (numpy) on the host, same inputs.  Prints one JSON object; not part of bench.py's contract line.
Usage: python tools/bench_aux.py [--frames 512] > profiles/r1_aux_bench.json"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import inputs_oracle as IO, metrics_oracle as MO      # CPU baseline leg only
from plr2_b200 import inputs, metrics

H, W = 480, 640
CAM = (312.9869, 241.3109, 1066.778, 1067.487)


def gpu_ms(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=512)
    ap.add_argument("--cpu-frames", type=int, default=16)
    args = ap.parse_args()
    F, N, M = args.frames, 1000, 2300
    rng = np.random.default_rng(0)
    masks = np.zeros((F, H, W), dtype=np.uint8)
    bbox = np.zeros((F, 4), dtype=np.int32)
    for f in range(F):
        r0, c0 = int(rng.integers(0, H - 120)), int(rng.integers(0, W - 160))
        masks[f, r0:r0 + 120, c0:c0 + 160] = rng.random((120, 160)) < 0.7
        bbox[f] = (r0, r0 + 120, c0, c0 + 160)
    depth = rng.integers(3000, 20000, size=(F, H, W)).astype(np.uint16)
    cam = np.array([[*CAM, 10000.0]] * F, dtype=np.float32)
    model = ((rng.random((F, M, 3)) - 0.5) * 0.2).astype(np.float32)
    R = np.stack([np.linalg.qr(rng.normal(size=(3, 3)))[0] for _ in range(F)])
    t = rng.normal(size=(F, 3)) * 0.5
    dm, db = torch.from_numpy(masks).cuda(), torch.from_numpy(bbox).cuda()
    dd, dc = torch.from_numpy(depth.view(np.int16)).cuda(), torch.from_numpy(cam).cuda()
    dmodel, dR, dt = torch.from_numpy(model).cuda(), torch.from_numpy(R).cuda(), torch.from_numpy(t).cuda()
    choose, _ = inputs.choose_from_mask(dm, db, N, seed=1)
    pose_a = torch.cat([dR.reshape(F, 9), dt], dim=1).float()
    pose_b = pose_a.clone(); pose_b[:, 9:] += 0.01
    dis = torch.from_numpy((rng.random(20000) * 0.12).astype(np.float32)).cuda()

    out = {"device": torch.cuda.get_device_name(0), "frames": F, "num_points": N, "num_mesh": M, "image": [H, W]}
    ms = gpu_ms(lambda: inputs.choose_from_mask(dm, db, N, seed=1))
    crop_bytes = F * 120 * 160
    out["choose_from_mask"] = {"ms": ms, "frames_per_s": F / ms * 1e3, "crop_mask_GB_s": crop_bytes / ms / 1e6}
    ms = gpu_ms(lambda: inputs.pointcloud(dd, db, choose, dc))
    out["backproject"] = {"ms": ms, "frames_per_s": F / ms * 1e3, "algorithmic_GB_s": F * N * (8 + 2 + 12) / ms / 1e6}
    ms = gpu_ms(lambda: inputs.transform_target(dmodel, dR, dt))
    out["transform_target"] = {"ms": ms, "frames_per_s": F / ms * 1e3, "algorithmic_GB_s": F * M * 24 / ms / 1e6}
    ms = gpu_ms(lambda: metrics.pose_metrics(dmodel, pose_a, pose_b, True, False))
    out["ADD"] = {"ms": ms, "frames_per_s": F / ms * 1e3}
    ms = gpu_ms(lambda: metrics.pose_metrics(dmodel, pose_a, pose_b, False, True), iters=5)
    out["ADDS"] = {"ms": ms, "frames_per_s": F / ms * 1e3, "Gpairs_s": F * M * M / ms / 1e6,
                   "fp32_frac": 8 * F * M * M / ms / 1e9 / 74.45}
    ms = gpu_ms(lambda: metrics.auc(dis), iters=5)
    out["auc_20000"] = {"ms": ms, "Gcompares_s": 20000 * 20000 / ms / 1e6}

    # host legs (oracle = the reference's numpy statements), bounded sample
    cf = min(args.cpu_frames, F)
    t0 = time.perf_counter()
    for f in range(cf):
        cnt = int(masks[f][bbox[f][0]:bbox[f][1], bbox[f][2]:bbox[f][3]].sum())
        c_mask = None
        if cnt > N:
            c_mask = np.zeros(cnt, dtype=int); c_mask[:N] = 1; rng.shuffle(c_mask)
        ch, _ = IO.compute_choose(masks[f], bbox[f], N, c_mask)
    t1 = time.perf_counter()
    for f in range(cf):
        IO.pointcloud(depth[f], bbox[f], ch, *CAM, np.uint16(10000))
    t2 = time.perf_counter()
    for f in range(cf):
        IO.transform_target(model[f].astype(np.float64), R[f], t[f][None])
    t3 = time.perf_counter()
    T = lambda f: np.vstack([np.hstack([R[f], t[f][:, None]]), [0, 0, 0, 1]])
    for f in range(min(cf, 4)):
        MO.adds(model[f], T(f), T(f))
    t4 = time.perf_counter()
    MO.auc(dis.cpu().numpy())
    t5 = time.perf_counter()
    out["cpu_oracle"] = {"cores": 1, "sample_frames": cf,
                         "choose_frames_per_s": cf / (t1 - t0), "pointcloud_frames_per_s": cf / (t2 - t1),
                         "transform_target_frames_per_s": cf / (t3 - t2), "ADDS_frames_per_s": min(cf, 4) / (t4 - t3),
                         "auc_20000_ms": (t5 - t4) * 1e3,
                         "note": "pointcloud rebuilds xmap/ymap per call as the oracle is written; dataset.py builds them once"}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
