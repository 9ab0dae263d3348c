"""Config C5 of BASELINE.json: the data-parallel training step of the estimator stage.

Per optimiser step and rank: PoseNet (backbone output given) forward on `frames_per_gpu` frames (the reference's
`accumulate_grad_batches: 8`, yaml/exp/exp_load_and_fit.yml:13; tools/train.py:32,152-155) -> `Loss` with
num_pt_mesh_large = 2300 (ADD-S for symmetric classes: N*M^2 = 5.3e9 nearest-neighbour pairs per such frame) ->
backward -> ONE gradient all-reduce (NCCL over NVLink; the only collective of the whole path) -> Adam.
Gradients are AVERAGED over the frames of the global batch (Lightning's accumulation semantics, tools/lightning.py).

Frames are independent, so any assignment of the global batch to ranks is a valid data-parallel step.  The cost of a
frame is dominated by whether its class is symmetric (0.75 ms of ADD-S search against ~0.05 ms of ADD), so the
global batch is dealt cost-aware (`parallel.deal_by_cost`) instead of in contiguous shards: no rank waits in the
all-reduce for a neighbour that happened to draw more symmetric frames.

The whole step (forward, loss, backward, all-reduce, fused Adam: ~300 launches for 8 frames) can be captured in one
CUDA graph (`cuda_graph=True`): inputs are static device buffers, the optimiser runs in capturable mode.
"""
import numpy as np
import torch
import torch.distributed as dist

from . import ops, parallel, synth
from .lib.loss import Loss
from .lib.network import PoseNet


def synthetic_global_batch(shape, n_frames, seed=5000, sym_every=4):
    """Frames of the global batch: one in `sym_every` has a symmetric class (YCB: 5 of 21 classes are symmetric),
    classes cycle through the respective lists, order shuffled like a data loader would deliver it."""
    sym = list(shape.sym_list)
    non = [c for c in range(shape.num_obj) if c not in sym]
    rng = np.random.Generator(np.random.PCG64(seed))
    objs = [sym[(i // sym_every) % len(sym)] if i % sym_every == 1 else non[i % len(non)] for i in range(n_frames)]
    order = rng.permutation(n_frames)
    return [synth.make_frame(shape, seed + 1 + int(i), obj=objs[int(i)]) for i in order]


class C5TrainStep:
    def __init__(self, device, world=1, rank=0, frames_per_gpu=8, precision=ops.PREC_BF16X2_ALL,
                 knn_mode=ops.KNN_LOSS_DEFAULT, overlap=False, cuda_graph=False, shape=synth.YCB_REFINE, lr=1e-4,
                 deal="cost", seed=5000):
        self.dev, self.world, self.rank = torch.device(device), world, rank
        self.shape, self.frames_per_gpu = shape, frames_per_gpu
        self.precision, self.knn_mode = precision, knn_mode
        net = PoseNet(shape.num_points, shape.num_obj, precision=precision)
        net.cnn = torch.nn.Identity()
        net.load_state_dict({k: torch.from_numpy(v) for k, v in synth.make_params("posenet", shape.num_obj, 1).items()},
                            strict=False)
        self.net = net.to(self.dev)
        self.crit = Loss(shape.num_mesh, list(shape.sym_list), knn_mode)
        self.params = self.net.hot_path_parameters()
        self.bucket = parallel.GradBucket(self.params)
        self.net.accumulate_grads_in_place(True)      # the bucket's views are the .grad tensors: one kernel adds into them
        self.in_op = parallel.InOpGradSync(average=True).attach(self.net) if (overlap and world > 1) else None
        self.cuda_graph = bool(cuda_graph)
        self.opt = torch.optim.Adam(self.params, lr=lr, fused=True, capturable=self.cuda_graph)
        # the global batch is the same on every rank (seeded); each rank keeps its share
        frames = synthetic_global_batch(shape, frames_per_gpu * world, seed)
        cost = [parallel.frame_cost(int(f["idx"]), shape.sym_list, shape.num_points, shape.num_mesh) for f in frames]
        if deal == "cost":
            mine = parallel.deal_by_cost(cost, world)[rank]
        else:
            a, b = parallel.shard_range(len(frames), world, rank)
            mine = list(range(a, b))
        self.frame_ids = mine
        self.sym_frames = int(sum(int(frames[i]["idx"]) in shape.sym_list for i in mine))
        frs = [frames[i] for i in mine]
        st = lambda k: torch.from_numpy(np.stack([f[k] for f in frs])).to(self.dev)
        self.out_img, self.points = st("out_img"), st("points")
        self.choose, self.obj = st("choose")[:, None], st("idx")[:, None]
        self.target, self.model = st("target"), st("model_points")
        self.w = 0.015
        self._graph = None
        self._loss = None
        self.graph_error = None

    # -- one optimiser step, eager ----------------------------------------------------------------
    def _step(self, ar_events=None):
        self.bucket.zero()
        pr, pt, pc, _ = self.net(self.out_img, self.points, self.choose, self.obj)
        loss, dis, _, _ = self.crit(pr, pt, pc, self.target, self.model, self.obj, self.points, self.w, False)
        loss.mean().backward()                       # average over the rank's frames
        if ar_events is not None:
            ar_events[0].record()
        if self.in_op is None:
            self.bucket.all_reduce(average=True)     # then over the ranks (in_op: already done inside the backward)
        if ar_events is not None:
            ar_events[1].record()
        self.opt.step()
        return loss

    def capture(self, warmup=3):
        """Capture the step in a CUDA graph (after `warmup` eager steps on a side stream).  Returns True on success;
        on failure the step stays eager and `graph_error` says why."""
        try:
            cur = torch.cuda.current_stream(self.dev)
            side = torch.cuda.Stream(device=self.dev)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                for _ in range(warmup):
                    self._step()
            cur.wait_stream(side)
            torch.cuda.synchronize(self.dev)
            if self.world > 1:
                dist.barrier()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                loss = self._step()
            self._graph, self._loss = g, loss
            return True
        except Exception as exc:                      # noqa: BLE001 - reported, never hidden
            self.graph_error = f"{type(exc).__name__}: {exc}"[:300]
            self._graph = None
            return False

    def step(self, ar_events=None):
        if self._graph is not None:
            self._graph.replay()
            return self._loss
        return self._step(ar_events)

    # -- timing ---------------------------------------------------------------------------------
    def run(self, steps=10, warmup=3):
        """`warmup` untimed steps, then `steps` timed ones (CUDA events, max over ranks, no host synchronisation
        inside the timed region).  Returns the C5 record."""
        dev, world = self.dev, self.world
        if self.cuda_graph and self._graph is None and self.graph_error is None:
            self.capture(max(3, warmup))
        for _ in range(max(warmup, 3)):
            self.step()
        graphed = self._graph is not None
        ar_ev = None if graphed else [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(steps)]
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for it in range(steps):
            loss = self.step(None if graphed else ar_ev[it])
        e1.record()
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1)
        ar_ms = None if graphed else sum(a.elapsed_time(b) for a, b in ar_ev)
        chk = torch.stack([p.detach().double().sum() for p in self.params]).sum().reshape(1)
        sym = [self.sym_frames]
        if world > 1:
            t = torch.tensor([ms, ar_ms if ar_ms is not None else 0.0], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t[0])
            ar_ms = float(t[1]) if ar_ms is not None else None
            lo, hi = chk.clone(), chk.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            consistent = bool((hi - lo).abs().item() <= 1e-9 * abs(hi.item()))
            s = torch.zeros(world, dtype=torch.int64, device=dev)
            s[self.rank] = self.sym_frames
            dist.all_reduce(s)
            sym = [int(v) for v in s.tolist()]
        else:
            consistent = True
        shp = self.shape
        return {
            "config": "C5: PoseNet fwd+bwd + Loss (ADD-S for symmetric classes, M=%d), %d frames/GPU, one gradient "
                      "all-reduce (average) + fused Adam" % (shp.num_mesh, self.frames_per_gpu),
            "n_gpus": world, "steps": steps, "ms_per_step": ms / steps,
            "frames_per_s": world * self.frames_per_gpu * steps / (ms / 1e3),
            "precision": {ops.PREC_FP32: "fp32 SIMT", ops.PREC_BF16X2: "fp32 forward, tensor-core backward (PLR2_PREC_BF16X2)",
                          ops.PREC_BF16X2_ALL: "tensor-core forward (sign-safe ReLU) and backward (PLR2_PREC_BF16X2_ALL)"}[self.precision],
            "knn_mode": "fast" if self.knn_mode == ops.KNN_FAST else "exact",
            "cuda_graph": graphed, "cuda_graph_error": self.graph_error,
            "allreduce_ms_per_step": (ar_ms / steps) if ar_ms is not None else None,
            "grad_bucket_mb": self.bucket.numel * 4 / 1e6,
            "grad_delivery": "one accumulating scatter onto the bucket's .grad views (accumulate_grads_in_place)",
            "allreduce": ("inside the backward op, head weights overlapped with the trunk backward" if self.in_op
                          else "one flat bucket after the backward") if world > 1 else "none (1 GPU)",
            "sharding": "cost-aware deal of the global batch (symmetric frames round-robin)",
            "sym_frames_per_rank": sym,
            "replicas_consistent": consistent, "loss_finite": bool(torch.isfinite(loss).all()),
        }
