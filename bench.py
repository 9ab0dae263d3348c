#!/usr/bin/env python
"""Benchmark of the DenseFusion dense-fusion + loss hot path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W              # this repo (CUDA, sm_100a)
    python bench.py --impl reference --gpus N --steps K ...    # CPU restatement of the reference's path

Workload (config C4 of BASELINE.json / SURVEY.md §8d): per step, FRAMES = 512 synthetic YCB-shaped frames
(N=1000 points, 21 classes, 5 of them symmetric, 120x120 crop, M=2300 mesh points) go through PoseNet
(backbone output given) -> Loss(refine) -> 2 x (PoseRefineNet -> Loss_refine with ADD-S for symmetric
classes).  One pose per frame => metric poses/sec.  Frames are independent: rank r owns the contiguous shard
shard_range(512, world, r) ("strong" scaling, the 512/G split BASELINE states); no data-path collective.  The
weak-scaling curve (512 frames PER GPU) is measured in the same run and reported under `weak`.
The line also carries the other BASELINE configs: `c2` (PoseNet + ADD-S Loss, M=500), `c3_bs1_latency_ms`
(one frame, plain launches and CUDA-graph replay) and `train_c5` (data-parallel training step with the NCCL
gradient all-reduce).  One JSON line is printed by rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from plr2_b200 import synth  # noqa: E402

METRIC = "poses/sec (PoseNet+2 refine+ADD-S)"
SHAPE = synth.YCB_REFINE           # N=1000, O=21, M=2300, 120x120
DISTINCT = 32                      # distinct synthetic frames generated; tiled up to FRAMES


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "sm_max_mhz": d.get("sm_max_mhz", 1965.0), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "sm_max_mhz": 1965.0,
            "source": "fallback"}


def make_workload(frames, seed0):
    """dict of CPU tensors (frames, ...) in the reference's per-frame layout."""
    base = [synth.make_frame(SHAPE, seed0 + i, choose_mode="mask") for i in range(min(DISTINCT, frames))]
    pick = [base[i % len(base)] for i in range(frames)]
    st = lambda k: torch.from_numpy(np.stack([f[k] for f in pick]))
    return {"out_img": st("out_img"), "points": st("points"), "choose": st("choose")[:, None, :],
            "obj": st("idx")[:, None], "target": st("target"), "model_points": st("model_points")}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            if not (t0 - 0.05 <= ts <= t1 + 0.15):
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except Exception:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------------------
# CPU restatement of the reference's path (oracle port) — the reference arm and the cpu_baseline leg
def cpu_reference_step(work, params_pn, params_rf, frames):
    from oracle import densefusion_oracle as DO      # the ONLY place bench.py touches oracle/
    out = []
    with torch.no_grad():
        for f in range(frames):
            obj = int(work["obj"][f, 0])
            pr, pt, pc, emb = DO.posenet_forward(params_pn, work["out_img"][f], work["points"][f],
                                                 work["choose"][f, 0], obj, SHAPE.num_obj)
            _, dis, npts, ntgt, _ = DO.pose_loss(pr, pt, pc, work["target"][f], work["model_points"][f], obj,
                                                 work["points"][f], 0.015, True, SHAPE.sym_list)
            for _ in range(2):
                rr, rt = DO.refiner_forward(params_rf, npts, emb, obj, SHAPE.num_obj)
                dis, npts, ntgt, _ = DO.refine_loss(rr, rt, ntgt, work["model_points"][f], obj, npts, SHAPE.sym_list)
            out.append(float(dis))
    return out


def time_cpu_reference(work, params_pn, params_rf, frames, steps, warmup):
    for _ in range(warmup):
        cpu_reference_step(work, params_pn, params_rf, min(frames, 2))
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_reference_step(work, params_pn, params_rf, frames)
    dt = time.perf_counter() - t0
    return frames * steps / dt, dt


def time_gpu_eager(work, pn, rf, frames, dev):
    """Same-box GPU baseline: the reference's algorithm as eager PyTorch fp32 ops on the B200
    (TF32 off), one frame at a time like tools/train.py:182-188.  Uses the oracle's restatement of the
    network/loss (the reference's modules cannot travel) with the ADD-S nearest neighbour as torch.cdist +
    argmin (the reference's knn.cu needs THC and cannot be built).  Reported for context only."""
    from oracle import densefusion_oracle as DO
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    saved = DO.nearest_index
    DO.nearest_index = lambda ref, qry, contract=True: torch.cdist(qry, ref).argmin(dim=1)
    try:
        p = {k: torch.from_numpy(v).to(dev) for k, v in pn.items()}
        r = {k: torch.from_numpy(v).to(dev) for k, v in rf.items()}
        w = {k: v[:frames].to(dev) for k, v in work.items()}

        def run(n):
            out = []
            with torch.no_grad():
                for f in range(n):
                    obj = int(w["obj"][f, 0])                         # host sync, like idx[0].item() (lib/loss.py:40)
                    pr, pt, pc, emb = DO.posenet_forward(p, w["out_img"][f], w["points"][f], w["choose"][f, 0], obj, SHAPE.num_obj)
                    _, dis, npts, ntgt, _ = DO.pose_loss(pr, pt, pc, w["target"][f], w["model_points"][f], obj, w["points"][f],
                                                         0.015, True, SHAPE.sym_list)
                    for _ in range(2):
                        rr, rt = DO.refiner_forward(r, npts, emb, obj, SHAPE.num_obj)
                        dis, npts, ntgt, _ = DO.refine_loss(rr, rt, ntgt, w["model_points"][f], obj, npts, SHAPE.sym_list)
                    out.append(float(dis))
            return out
        run(2)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        run(frames)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    finally:
        DO.nearest_index = saved
    return {"value": frames / dt, "unit": "poses/s", "kind": "eager PyTorch fp32 (TF32 off) on the same B200, bs=1 loop",
            "sample": f"{frames} frames ({dt:.2f} s)"}


def knn_roofline(pk, ref_kernel=True):
    """ADD-S 1-NN kernel at config C2 size (500 refs x 500 000 queries per frame, 8 frames), both ranking modes, and - as the
    same-box baseline SURVEY.md 2 names - the reference's OWN CUDA kernels (lib/knn/src/cuda/knn.cu compiled unmodified for
    sm_100a into oracle/_ref/libknn_cuda_ref.so; 1 GB distance matrix per frame, driven like lib/knn/src/knn.h:29-41)."""
    from plr2_b200.lib.knn import knn_indices
    g = torch.Generator(device="cuda").manual_seed(0)
    ref = torch.randn(8, 3, 500, device="cuda", generator=g) * 0.05
    qry = torch.randn(8, 3, 500_000, device="cuda", generator=g) * 0.05
    peak_tflops = 148 * 128 * 2 * pk["sm_max_mhz"] * 1e6 / 1e12
    pairs = 8 * 500 * 500_000
    out = {}
    for name, mode in (("exact", 0), ("fast", 1)):
        for _ in range(3):
            knn_indices(ref, qry, 1, mode)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            knn_indices(ref, qry, 1, mode)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        out[name] = {"Gpairs_s": pairs / ms / 1e6, "achieved": 8 * pairs / ms / 1e9, "peak": peak_tflops,
                     "unit": "TFLOP/s", "frac": 8 * pairs / ms / 1e9 / peak_tflops, "bound": "fp32",
                     "ms_per_launch": ms, "pairs_per_launch": pairs}
    if ref_kernel:
        try:
            from oracle import knn_oracle as KO          # baseline leg: the reference's kernel is the thing compared WITH
            if not KO.have_reference_cuda_knn():
                raise FileNotFoundError("oracle/_ref/libknn_cuda_ref.so not built")
            scratch = torch.empty(500 * 500_000, dtype=torch.float32, device="cuda")
            ridx = KO.knn_reference_cuda(ref, qry, 1, scratch=scratch)
            same = bool(torch.equal(ridx, knn_indices(ref, qry, 1, 0)))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(2):
                KO.knn_reference_cuda(ref, qry, 1, scratch=scratch)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 2
            out["reference_cuda_kernel"] = {
                "Gpairs_s": pairs / ms / 1e6, "ms": ms, "pairs": pairs,
                "kind": "lib/knn/src/cuda/knn.cu (cuComputeDistanceGlobal + cuInsertionSort) unmodified, sm_100a, scratch held",
                "indices_equal_to_exact_mode": same,
                "speedup_exact": out["exact"]["Gpairs_s"] / (pairs / ms / 1e6),
                "speedup_fast": out["fast"]["Gpairs_s"] / (pairs / ms / 1e6)}
            del scratch
        except Exception as exc:                      # noqa: BLE001 - reported, not hidden
            out["reference_cuda_kernel"] = {"unavailable": f"{type(exc).__name__}: {exc}"[:200]}
    return out


def _ev_time(fn, reps, warm=2, dev=None):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1) / reps


def measure_c2(pipe, dev, knn_mode):
    """BASELINE config 2: YCB-shaped PoseNet (backbone output given) + ADD-S `Loss` on symmetric class 12, N = 1000,
    num_pt_mesh_small = 500 (2.5e8 nearest-neighbour pairs per frame), one B200.  Inference (64 frames per launch chain)
    and training (8 frames, forward + backward of network and loss)."""
    from plr2_b200.lib.loss import Loss
    shp = synth.YCB
    crit = Loss(shp.num_mesh, list(shp.sym_list), knn_mode)
    base = [synth.make_frame(shp, 2000 + i, obj=12, choose_mode="mask") for i in range(8)]

    def batch(n):
        st = lambda k: torch.from_numpy(np.stack([base[i % 8][k] for i in range(n)])).to(dev)
        return st("out_img"), st("points"), st("choose")[:, None], st("idx")[:, None], st("target"), st("model_points")
    net = pipe.estimator
    img, pts, ch, ob, tg, md = batch(64)

    def fwd():
        with torch.no_grad():
            pr, pt, pc, _ = net.forward_features(img, pts, ch, ob)
            return crit(pr, pt, pc, tg, md, ob, pts, 0.015, False)
    ms = _ev_time(fwd, 10, dev=dev)
    img8, pts8, ch8, ob8, tg8, md8 = batch(8)
    params = net.hot_path_parameters()

    def train():
        for p in params:
            p.grad = None
        pr, pt, pc, _ = net.forward_features(img8, pts8, ch8, ob8)
        loss, _, _, _ = crit(pr, pt, pc, tg8, md8, ob8, pts8, 0.015, False)
        loss.mean().backward()
    net.train()
    ms_t = _ev_time(train, 10, dev=dev)
    net.eval()
    for p in params:
        p.grad = None
    return {"config": "C2: PoseNet + Loss (ADD-S, class 12), N=1000, M=500", "inference_frames_per_s": 64 / (ms / 1e3),
            "inference_ms_per_64_frames": ms, "adds_gpairs_per_s_inside_step": 64 * 2.5e8 / (ms / 1e3) / 1e9,
            "train_fwd_bwd_frames_per_s": 8 / (ms_t / 1e3), "train_ms_per_8_frames": ms_t}


def measure_c3_latency(pn, rf, dev, knn_mode):
    """BASELINE config 3: ONE YCB frame through PoseNet -> Loss(refine) -> 2 x (PoseRefineNet -> Loss_refine), bs = 1 -
    what a tools/lightning.py / tools/train.py:182-188 user waits for per frame.  ~32 launches: launch-bound, so the
    chain is also measured as a CUDA-graph replay.  Wall time per call including the final synchronisation."""
    from plr2_b200 import ops
    from plr2_b200.pipeline import DenseFusionPipeline, INPUT_KEYS
    out = {}
    for label, obj in (("nonsym", 3), ("sym", 12)):
        fr = synth.make_frame(SHAPE, 4242 + obj, obj=obj, choose_mode="mask")
        t = lambda k: torch.from_numpy(fr[k])[None].to(dev)
        work = {"out_img": t("out_img"), "points": t("points"), "choose": t("choose")[:, None], "obj": t("idx").reshape(1, 1),
                "target": t("target"), "model_points": t("model_points")}
        rec = {}
        for mode, graph in (("plain", False), ("cuda_graph", True)):
            pipe = DenseFusionPipeline(SHAPE.num_points, SHAPE.num_obj, SHAPE.num_mesh, SHAPE.sym_list, device=dev,
                                       precision=ops.PREC_BF16X2, knn_mode=knn_mode, chunk_frames=1, cuda_graph=graph)
            pipe.load_params(pn, rf)
            for _ in range(5):
                res = pipe.infer(*(work[k] for k in INPUT_KEYS))
            torch.cuda.synchronize(dev)
            n = 200
            t0 = time.perf_counter()
            for _ in range(n):
                res = pipe.infer(*(work[k] for k in INPUT_KEYS))
                torch.cuda.synchronize(dev)
            rec[mode + "_ms"] = (time.perf_counter() - t0) / n * 1e3
            rec[mode + "_dis"] = float(res[0])
            del pipe
        out[label] = rec
    return out


def measure_train_c5(args, world, rank, dev):
    """BASELINE config 5 under the driver's clock: data-parallel training step, 8 frames per GPU, M = 2300, one NCCL
    gradient all-reduce per step (plr2_b200/train.py).  Primary record: all contractions on tensor cores, whole step
    replayed as one CUDA graph; the eager variants are listed beside it."""
    from plr2_b200 import ops
    from plr2_b200.train import C5TrainStep
    knn_mode = ops.KNN_FAST if args.knn == "fast" else ops.KNN_EXACT
    recs = {}
    for name, prec, graph in (("tc_all_graph", ops.PREC_BF16X2_ALL, True), ("tc_all_eager", ops.PREC_BF16X2_ALL, False),
                              ("fp32fwd_tcbwd_eager", ops.PREC_BF16X2, False)):
        try:
            ts = C5TrainStep(dev, world, rank, 8, precision=prec, knn_mode=knn_mode, cuda_graph=graph)
            recs[name] = ts.run(args.train_steps, 3)
            del ts
        except Exception as exc:                      # noqa: BLE001
            recs[name] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
        torch.cuda.empty_cache()
    ok = [k for k in ("tc_all_graph", "tc_all_eager", "fp32fwd_tcbwd_eager") if "frames_per_s" in recs[k]]
    best = dict(recs[ok[0]]) if ok else {"error": "no variant ran"}
    best["variant"] = ok[0] if ok else None
    best["variants"] = {k: ({kk: v[kk] for kk in ("ms_per_step", "frames_per_s", "allreduce_ms_per_step", "cuda_graph",
                                                  "cuda_graph_error", "replicas_consistent") if kk in v} if "frames_per_s" in v else v)
                        for k, v in recs.items()}
    return best


# --------------------------------------------------------------------------------------------------
def dist_setup(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        backend = "gloo" if args.impl == "reference" else "nccl"
        if args.impl != "reference":
            torch.cuda.set_device(local)
        dist.init_process_group(backend=backend)
    return world, rank, local


def run_reference(args, world, rank):
    """Reference arm: the reference's CPU implementation of the path (oracle port; the reference's own
    Python modules cannot travel to the GPU box) on the box's host cores, rank 0 only."""
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    frames = args.ref_frames
    work = make_workload(frames, 1000)
    pn = {k: torch.from_numpy(v) for k, v in synth.make_params("posenet", SHAPE.num_obj, 1).items()}
    rf = {k: torch.from_numpy(v) for k, v in synth.make_params("refiner", SHAPE.num_obj, 2).items()}
    rate, dt = time_cpu_reference(work, pn, rf, frames, args.steps, min(args.warmup, 3))
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": "poses/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C4: 512 YCB-shaped frames per step sharded 512/G over the GPUs, PoseNet(out_img given)+Loss(refine)+"
                               "2x(PoseRefineNet+Loss_refine ADD-S)",
                   "frames_per_step": frames, "sample_of_frames_per_step": 512, "num_points": SHAPE.num_points, "num_mesh": SHAPE.num_mesh,
                   "num_obj": SHAPE.num_obj},
        "cpu_baseline": {"value": rate, "unit": "poses/s", "cores": torch.get_num_threads(), "kind": "port",
                         "sample": f"{frames} frames/step x {args.steps} steps (oracle: torch CPU fp32 + C/OpenMP 1-NN)"},
        "e2e": {"value": rate, "unit": "poses/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def time_with_backbone(pipe, devw, frames, dev, chunk=64):
    """SURVEY 8d asks for the metric with and without the (out-of-scope) PSPNet/ResNet-18 backbone: the same
    device-resident chain, but `out_img` is produced by the backbone (PyTorch, random-init weights) from synthetic
    `img` = randn(3,H,W) (lib/network.py:96, lib/pspnet.py:40-77).  `value` is the parity-preserving fp32 figure
    (TF32 off): the best of the fp32 variants of the batched runner of SURVEY 8f-1 (plr2_b200/backbone.py); the
    reduced-precision modes are listed beside it."""
    from plr2_b200.backbone import BatchedBackbone
    from plr2_b200.lib.network import ModifiedResnet
    h, w = devw["out_img"].shape[-2:]
    gen = torch.Generator(device=dev).manual_seed(5)
    img = torch.randn(frames, 3, h, w, device=dev, generator=gen)
    keys = [k for k in ("points", "choose", "obj", "target", "model_points")]
    # the deployment shape of the END-TO-END call (tools/lightning.py:115-116 hands PoseNet the RGB crop): every input
    # starts in pinned host memory - img (173 KB / frame) instead of the backbone output (1.84 MB / frame) - and the
    # per-frame distance comes back to the host
    himg = img.cpu().pin_memory()
    hsmall = {k: devw[k][:frames].cpu().pin_memory() for k in keys}
    e2e_bytes = (himg[0].numel() * 4 + sum(hsmall[k][0].numel() * hsmall[k].element_size() for k in keys)) * frames

    def measure(make_out_img, reps=3, from_host=False):
        @torch.no_grad()
        def step():
            out = None
            for a in range(0, frames, chunk):
                b = min(frames, a + chunk)
                if from_host:
                    im = himg[a:b].to(dev, non_blocking=True)
                    small = [hsmall[k][a:b].to(dev, non_blocking=True) for k in keys]
                    out = pipe.infer(make_out_img(im), *small).to("cpu", non_blocking=True)
                else:
                    out = pipe.infer(make_out_img(img[a:b]), *(devw[k][a:b] for k in keys))
            return out
        for _ in range(2):
            res = step()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            res = step()
        e1.record()
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1) / reps
        return {"poses_per_s": frames / (ms / 1e3), "ms_per_step": ms, "results_finite": bool(torch.isfinite(res).all())}

    variants = {}
    for name, kw in (("eager_fp32_nchw", dict(precision="fp32", channels_last=False, cuda_graph=False)),
                     ("graph_fp32_nchw", dict(precision="fp32", channels_last=False, cuda_graph=True)),
                     ("graph_fp32_channels_last", dict(precision="fp32", channels_last=True, cuda_graph=True)),
                     ("graph_tf32_channels_last", dict(precision="tf32", channels_last=True, cuda_graph=True)),
                     ("graph_bf16_channels_last", dict(precision="bf16", channels_last=True, cuda_graph=True))):
        torch.manual_seed(0)
        bb = BatchedBackbone(ModifiedResnet().to(dev), batch=chunk, **kw)
        try:
            variants[name] = measure(bb)
            if name in ("graph_fp32_nchw", "graph_tf32_channels_last", "graph_bf16_channels_last"):
                variants[name]["e2e_from_img_poses_per_s"] = measure(bb, from_host=True)["poses_per_s"]
        except Exception as exc:                      # a variant that cannot run here is reported, not hidden
            variants[name] = {"error": f"{type(exc).__name__}: {exc}"[:200]}
        del bb
        torch.cuda.empty_cache()
    fp32 = [v for k, v in variants.items() if "fp32" in k and "poses_per_s" in v]
    best = max(fp32, key=lambda v: v["poses_per_s"]) if fp32 else {}
    return {"value": best.get("poses_per_s"), "unit": "poses/s", "frames_per_step": frames,
            "e2e_from_img": {"h2d_bytes_per_step": e2e_bytes, "d2h_bytes_per_step": frames * 4,
                             "note": "inputs (img, points, choose, obj, target, model_points) start in pinned host memory, the PyTorch "
                                     "backbone runs on the device, the per-frame distance returns to the host",
                             "poses_per_s": {k: v.get("e2e_from_img_poses_per_s") for k, v in variants.items()
                                             if isinstance(v, dict) and "e2e_from_img_poses_per_s" in v}},
            "backbone": f"PSPNet/ResNet-18 in PyTorch, fp32 (TF32 off), batches of {chunk} frames, random-init weights; "
                        "tf32/bf16 variants are reduced precision (not parity modes)",
            "variants": variants}


def run_ours(args, world, rank, local):
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this path has no CPU fallback; use --impl reference)")
    from plr2_b200 import _cabi, ops
    from plr2_b200.parallel import bind_to_gpu_numa, shard_range
    from plr2_b200.pipeline import DenseFusionPipeline, INPUT_KEYS
    import ctypes as C
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    L = _cabi.lib()
    pk = peaks()
    total = args.frames                                  # C4: 512 frames in all, sharded 512 / G
    lo, hi = shard_range(total, world, rank)
    n_local = hi - lo
    knn_mode = ops.KNN_FAST if args.knn == "fast" else ops.KNN_EXACT
    # Each chunk's launch chain (~32 kernels of this library + the PyTorch glue between them) is replayed as a CUDA graph:
    # small shards are launch-bound outright, and even a 512-frame chunk gains ~5 % (no launch gaps).  The per-kernel CUDA
    # events of the roofline cannot live inside a replay, so they bracket the same launches in an eager pass of the same
    # steps straight after the timed region.
    use_graph = args.cuda_graph != "off"
    host_chunk = args.host_chunk if args.host_chunk > 0 else min(64, max(16, (n_local + 3) // 4))
    pipe = DenseFusionPipeline(SHAPE.num_points, SHAPE.num_obj, SHAPE.num_mesh, SHAPE.sym_list, device=dev,
                               precision=ops.PREC_BF16X2, knn_mode=knn_mode, chunk_frames=args.chunk,
                               host_chunk_frames=host_chunk, cuda_graph=False, max_graphs=16, copy_streams=args.copy_streams)
    pn = synth.make_params("posenet", SHAPE.num_obj, 1)
    rf = synth.make_params("refiner", SHAPE.num_obj, 2)
    pipe.load_params(pn, rf)
    numa_cpus = bind_to_gpu_numa(local) if world > 1 else None     # pinned buffers on the GPU's own node
    work_all = make_workload(total, 1000)                # the same global batch on every rank
    work = {k: v[lo:hi].contiguous() for k, v in work_all.items()}
    if args.host_alloc == "wc":                       # write-combined staging buffers (experiment for host-bound N = 8)
        from plr2_b200.pipeline import host_staging_buffer
        host = {}
        for k, v in work.items():
            host[k] = host_staging_buffer(tuple(v.shape), v.dtype, write_combined=True)
            host[k].copy_(v)
    else:
        host = {k: v.pin_memory() for k, v in work.items()}
    frame_bytes = DenseFusionPipeline.bytes_per_frame(host)
    # Device-resident input sets, used in rotation: consecutive steps never read inputs the previous step left in
    # the 126 MB L2 (one set when the shard alone is larger than twice the L2).
    l2_bytes = 126 << 20
    n_sets = 1 if frame_bytes * n_local >= 2 * l2_bytes else min(4, 1 + (2 * l2_bytes) // max(1, frame_bytes * n_local))
    devsets = [{k: v.to(dev, non_blocking=True).clone() for k, v in host.items()} for _ in range(n_sets)]
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world > 1:
            import torch.distributed as dist
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return ms

    def step(i, p=pipe, sets=devsets):
        d = sets[i % len(sets)]
        return p.infer(*(d[k] for k in INPUT_KEYS), return_pose=True)

    # ---- one eager step: launch count per step, and the result the parity check uses ------------
    step(0)
    torch.cuda.synchronize()
    l0 = L.plr2_launch_count()
    res, poses = step(0)
    torch.cuda.synchronize()
    launches_per_step = int(L.plr2_launch_count() - l0)
    res_eager = res.clone()
    pipe.cuda_graph = use_graph

    # ---- device-resident (value) ---------------------------------------------------------------
    for i in range(max(args.warmup, 3) + n_sets):
        step(i)
    barrier()
    if args.profile_step:
        # for `ncu --profile-from-start off`: exactly one device-resident step inside the capture range
        pipe.cuda_graph = False
        torch.cuda.profiler.start()
        step(0)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        return
    sampler = ClockSampler(local) if rank == 0 else None
    if not use_graph:
        L.plr2_profile_begin()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record()
    for i in range(args.steps):
        res, poses = step(i)
    e1.record()
    barrier()
    t1 = time.time()
    ms_dev = max_over_ranks(e0.elapsed_time(e1))
    tc_ms, tc_n, tc_fl = C.c_double(), C.c_longlong(), C.c_double()
    roofline_pass = "timed region"
    if use_graph:
        # the timed region replayed CUDA graphs (no per-kernel events inside a replay): bracket the same launches in an
        # eager pass of the same steps right behind it
        pipe.cuda_graph = False
        L.plr2_profile_begin()
        for i in range(args.steps):
            step(i)
        pipe.cuda_graph = True
        roofline_pass = "separate eager pass of the same steps (the timed region replays CUDA graphs)"
    L.plr2_profile_end(C.byref(tc_ms), C.byref(tc_n), C.byref(tc_fl))
    tc_exec = float(L.plr2_profile_executed_flops())
    clocks = sampler.stop(t0, t1) if sampler else None
    value = total * args.steps / (ms_dev / 1e3)
    res = res.clone()

    # ---- end to end from pinned host buffers (e2e) ---------------------------------------------
    for _ in range(3):
        pipe.infer_host(host)
    barrier()
    e0.record()
    for _ in range(args.steps):
        res_host = pipe.infer_host(host)
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    e2e = total * args.steps / (ms_e2e / 1e3)
    h2d = frame_bytes * n_local
    d2h = n_local * 4 + n_local * 12 * 4       # per-frame distance + refined pose (R | t)

    # ---- weak-scaling curve (N > 1): 512 frames PER GPU, plain launches ------------------------
    weak = None
    if world > 1 and args.weak_steps > 0:
        pipe.cuda_graph = False
        wk = make_workload(args.frames, 1000 + 7919 * rank)
        wdev = [{k: v.to(dev) for k, v in wk.items()}]
        for i in range(3):
            step(i, sets=wdev)
        barrier()
        e0.record()
        for i in range(args.weak_steps):
            step(i, sets=wdev)
        e1.record()
        barrier()
        ms_w = max_over_ranks(e0.elapsed_time(e1))
        weak = {"value": world * args.frames * args.weak_steps / (ms_w / 1e3), "unit": "poses/s", "scaling": "weak",
                "frames_per_gpu_per_step": args.frames, "steps": args.weak_steps, "ms_per_step": ms_w / args.weak_steps}
        del wdev, wk
        torch.cuda.empty_cache()
        pipe.cuda_graph = use_graph

    # ---- C5 training step (all ranks: it contains the gradient all-reduce) ----------------------
    train_c5 = measure_train_c5(args, world, rank, dev) if args.train_steps > 0 else None
    barrier()
    if rank != 0:
        return

    # ---- parity check of the timed results (rank 0) ---------------------------------------------
    rel = lambda a, b: float(((a - b).abs() / b.abs().clamp_min(1e-12)).max())
    parity = {"max_rel_err_device_path_vs_host_path": rel(res.cpu(), res_host),
              "max_rel_err_graph_vs_eager": rel(res, res_eager) if use_graph else None,
              "device_chunk_frames": min(args.chunk, n_local), "host_chunk_frames": host_chunk,
              "results_finite": bool(torch.isfinite(res_host).all()) and bool(torch.isfinite(res).all())}
    if args.check_frames > 0:
        torch.set_num_threads(os.cpu_count() or 1)
        pnt = {k: torch.from_numpy(v) for k, v in pn.items()}
        rft = {k: torch.from_numpy(v) for k, v in rf.items()}
        ids = sorted(set(int(i) for i in np.linspace(0, n_local - 1, args.check_frames)))
        sub = {k: v[ids] for k, v in work.items()}
        ref = torch.tensor(cpu_reference_step(sub, pnt, rft, len(ids)))     # the oracle is the checker, never the product
        got = res.cpu()[ids]
        parity.update({"frames_vs_cpu_oracle": ids, "max_rel_err_vs_cpu_oracle": rel(got, ref),
                       "sym_frames_checked": int(sum(int(work["obj"][i, 0]) in SHAPE.sym_list for i in ids))})
        parity["ok"] = bool(parity["max_rel_err_vs_cpu_oracle"] < 2e-3 and parity["max_rel_err_device_path_vs_host_path"] < 2e-3
                            and parity["results_finite"])
        if not parity["ok"]:
            print("bench.py: PARITY CHECK FAILED: %s" % json.dumps(parity), file=sys.stderr, flush=True)

    # ---- rooflines, CPU baseline (rank 0, N == 1 only for the CPU leg) --------------------------
    achieved = tc_fl.value / (tc_ms.value / 1e3) / 1e12 if tc_ms.value > 0 else 0.0
    executed = tc_exec / (tc_ms.value / 1e3) / 1e12 if tc_ms.value > 0 else 0.0
    peak = pk["bf16_tflops_sustained"]
    traffic, traffic_src = None, None   # dram__bytes_{read,write}.sum per launch of this kernel, from the committed ncu capture
    tp = os.path.join(ROOT, "profiles", "gemm_traffic.json")
    if os.path.exists(tp):
        tj = json.load(open(tp))
        traffic = tj.get("avg_dram_bytes_per_launch")
        traffic_src = "profiles/gemm_traffic.json (ncu --set full capture of a 512-frame step, not measured in this run)"
    roofline = {"bound": "tensor", "kernel": "gemm_bf16x2_kernel (tcgen05 cta_group::2 kind::f16 on split-bf16 operands: 3 MMAs per "
                                             "algorithmic MAC, 2 for the pooled-only layers)",
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic,
                "traffic_source": traffic_src,
                "executed_tflops": executed, "frac_executed": executed / peak,
                "launches": int(tc_n.value), "avg_launch_ms": tc_ms.value / max(1, tc_n.value),
                "share_of_step": tc_ms.value / (ms_dev if not use_graph else max(ms_dev, tc_ms.value)) if ms_dev else None,
                "measured_over": roofline_pass,
                "peak_source": pk["source"] + " bf16_tflops_sustained"}
    knn = knn_roofline(pk)
    cpu = None
    if world == 1 and args.cpu_frames > 0:
        torch.set_num_threads(os.cpu_count() or 1)
        pnt = {k: torch.from_numpy(v) for k, v in pn.items()}
        rft = {k: torch.from_numpy(v) for k, v in rf.items()}
        rate, dt = time_cpu_reference(work, pnt, rft, args.cpu_frames, 2, 1)
        cpu = {"value": rate, "unit": "poses/s", "cores": torch.get_num_threads(), "kind": "port",
               "sample": f"{args.cpu_frames} frames x 2 passes of the same workload ({dt:.1f} s; oracle: torch CPU fp32 + C/OpenMP 1-NN)"}
    c2 = c3 = None
    if world == 1 and not args.skip_configs:
        c2 = measure_c2(pipe, dev, knn_mode)
        c3 = measure_c3_latency(pn, rf, dev, knn_mode)
    graph_extra = None
    if world == 1 and not args.skip_configs:
        # the same device-resident step the other way round (eager launches when `value` replays graphs, and vice versa)
        pipe.cuda_graph = not use_graph
        ms_g = _ev_time(lambda: step(0), args.steps, warm=3, dev=dev)
        pipe.cuda_graph = use_graph
        graph_extra = {"value": total / (ms_g / 1e3), "ms_per_step": ms_g, "cuda_graph": not use_graph}
    with_bb = None
    if world == 1 and args.backbone_frames > 0:
        pipe.cuda_graph = False      # fresh input tensors every call here: chunk graphs (keyed on addresses) would be re-captured
        with_bb = time_with_backbone(pipe, devsets[0], min(args.backbone_frames, n_local), dev)
        pipe.cuda_graph = use_graph
    eager = None
    if world == 1 and args.eager_frames > 0:
        eager = time_gpu_eager(work, pn, rf, args.eager_frames, dev)
    line = {
        "metric": METRIC, "value": value, "unit": "poses/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16x2-split operands, f32 accumulate (f32 SIMT for small layers/losses)",
        "data": "synthetic",
        "config": {"workload": "C4: 512 YCB-shaped frames per step sharded 512/G over the GPUs, PoseNet(out_img given)+Loss(refine)+"
                               "2x(PoseRefineNet+Loss_refine ADD-S)",
                   "frames_per_step": total, "frames_per_gpu_per_step": n_local, "chunk_frames": args.chunk,
                   "host_chunk_frames": host_chunk, "cuda_graph": use_graph, "host_alloc": args.host_alloc, "copy_streams": args.copy_streams,
                   "numa_bound_cpus": len(numa_cpus) if numa_cpus else None, "num_points": SHAPE.num_points,
                   "num_mesh": SHAPE.num_mesh, "num_obj": SHAPE.num_obj, "crop": [SHAPE.height, SHAPE.width],
                   "choose": "pixels of a compact object mask (datasets/ycb/dataset.py:177-184)",
                   "knn_mode": args.knn,
                   "l2": "%d device-resident input set(s) of %.0f MB used in rotation (> 2x the 126 MB L2 between re-reads); "
                         "%.2f GB of activations stream through L2 per step" % (n_sets, frame_bytes * n_local / 1e6,
                                                                               n_local * 1000 * 13.5e3 / 1e9)},
        "e2e": {"value": e2e, "unit": "poses/s", "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": d2h * world,
                "ms_per_step": ms_e2e / args.steps, "h2d_gbs_per_rank": h2d / (ms_e2e / args.steps / 1e3) / 1e9},
        "gpu_launches": launches_per_step * args.steps,
        "gpu_launches_note": "%d kernels of this library per step and rank (counted on an eager step; graph replays launch the same kernels)" % launches_per_step,
        "clocks": clocks,
        "roofline": roofline,
        "parity_check": parity,
        "weak": weak,
        "train_c5": train_c5,
        "c2": c2,
        "c3_bs1_latency_ms": c3,
        "value_other_launch_mode": graph_extra,
        "roofline_knn": knn,
        "knn_gpairs_per_s": {k: v["Gpairs_s"] for k, v in knn.items() if "Gpairs_s" in v},
        "cpu_baseline": cpu,
        "gpu_eager_baseline": eager,
        "with_backbone": with_bb,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=512, help="frames per step over ALL GPUs (config C4: sharded 512/G)")
    ap.add_argument("--chunk", type=int, default=512, help="frames per launch chunk, device-resident path")
    ap.add_argument("--host-chunk", type=int, default=0,
                    help="frames per chunk of the host (e2e) path: H2D of chunk i+1 overlaps chunk i (0 = min(64, max(16, shard/4)))")
    ap.add_argument("--cuda-graph", default="on", choices=["auto", "on", "off"],
                    help="replay each chunk's launch chain as a CUDA graph (on / auto) or launch eagerly (off)")
    ap.add_argument("--copy-streams", type=int, default=1, help="streams the H2D copies of a host chunk are spread over (e2e path)")
    ap.add_argument("--host-alloc", default="pinned", choices=["pinned", "wc"],
                    help="host buffers of the e2e path: page-locked (default) or page-locked write-combined")
    ap.add_argument("--weak-steps", type=int, default=5, help="steps of the extra weak-scaling measurement at N > 1 (0 = skip)")
    ap.add_argument("--train-steps", type=int, default=10, help="steps of the C5 training record (0 = skip)")
    ap.add_argument("--check-frames", type=int, default=4, help="frames of the timed result compared with the CPU oracle (rank 0)")
    ap.add_argument("--skip-configs", action="store_true", help="skip the C2 / C3 / graph-replay side records")
    ap.add_argument("--knn", default="fast", choices=["exact", "fast"],
                    help="ranking arithmetic of the 1-NN scan fused into the ADD-S losses (ops.KNN_LOSS_DEFAULT = fast)")
    ap.add_argument("--cpu-frames", type=int, default=8, help="frames of the CPU baseline sample (0 = skip)")
    ap.add_argument("--profile-step", action="store_true",
                    help="run ONE step between cudaProfilerStart/Stop and exit (for ncu --profile-from-start off)")
    ap.add_argument("--backbone-frames", type=int, default=256,
                    help="frames of the extra with-backbone measurement (N=1 only; 0 disables)")
    ap.add_argument("--eager-frames", type=int, default=16,
                    help="frames of the same-box eager-PyTorch GPU baseline (reference algorithm, fp32, bs=1 loop; 0 = skip)")
    ap.add_argument("--ref-frames", type=int, default=4, help="frames per step of the reference arm")
    args = ap.parse_args()
    world, rank, local = dist_setup(args)
    if args.impl == "reference":
        run_reference(args, world, rank)
    else:
        run_ours(args, world, rank, local)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
