"""Frame-batched DenseFusion inference / evaluation pipeline — the call a user of this repo makes.

Reproduces, for F independent frames at once, the reference's per-frame test loop
(tools/train.py:182-188):

    pred_r, pred_t, pred_c, emb = estimator(img, points, choose, idx)
    _, dis, new_points, new_target = criterion(pred_r, pred_t, pred_c, target, model_points, idx, points, w, refine_start)
    for ite in range(iteration):                       # kernel (5): stays on the device, no host sync
        pred_r, pred_t = refiner(new_points, emb, idx)
        dis, new_points, new_target = criterion_refine(pred_r, pred_t, new_target, model_points, idx, new_points)

with the PSPNet backbone left out (its output `out_img` is the input here).  Frames are processed in
chunks so the activation workspace stays bounded; `infer_host` streams chunks from pinned host memory
on a copy stream while the previous chunk computes.
"""
import torch

from . import ops
from .lib.loss import Loss
from .lib.loss_refiner import Loss_refine
from .lib.network import PoseNet, PoseRefineNet

INPUT_KEYS = ("out_img", "points", "choose", "obj", "target", "model_points")


class DenseFusionPipeline:
    def __init__(self, num_points, num_obj, num_mesh, sym_list, device="cuda", precision=ops.PREC_BF16X2,
                 knn_mode=ops.KNN_LOSS_DEFAULT, refine_iters=2, chunk_frames=512, w=0.015, refine_start=True,
                 with_backbone=False, host_chunk_frames=64, cuda_graph=False, max_graphs=8, copy_streams=1):
        self.device = torch.device(device)
        self.estimator = PoseNet(num_points, num_obj, precision=precision)
        if not with_backbone:
            self.estimator.cnn = torch.nn.Identity()
        self.estimator = self.estimator.to(self.device).eval()
        self.refiner = PoseRefineNet(num_points, num_obj, precision=precision).to(self.device).eval()
        self.criterion = Loss(num_mesh, list(sym_list), knn_mode)
        self.criterion_refine = Loss_refine(num_mesh, list(sym_list), knn_mode)
        self.refine_iters = refine_iters
        self.chunk_frames = chunk_frames            # device-resident path: large chunks amortise launches and tile tails
        self.host_chunk_frames = host_chunk_frames  # host path: smaller chunks so the H2D copy of chunk i+1 overlaps chunk i
        self.w = w
        self.refine_start = refine_start
        self._staging = None
        self.copy_streams = copy_streams            # host path: H2D copies of a chunk spread over this many streams (out_img split by frames)
        # cuda_graph: the launch chain of a chunk (~32 kernels + the PyTorch glue around them) is captured once per
        # (input addresses, shapes, parameter versions) and replayed; pays off where a chunk is launch-bound
        # (bs = 1 latency, small shards of a strongly scaled batch).  Results are identical: same kernels, same order.
        self.cuda_graph = cuda_graph
        self.max_graphs = max_graphs
        self._graphs = {}

    def load_params(self, posenet_params, refiner_params):
        """dicts name -> numpy/torch array in the reference's state-dict naming."""
        as_t = lambda d: {k: torch.as_tensor(v) for k, v in d.items()}
        self.estimator.load_state_dict(as_t(posenet_params), strict=False)
        self.refiner.load_state_dict(as_t(refiner_params))

    # -- one chunk, everything already on the device -------------------------------------------
    def _chunk(self, out_img, points, choose, obj, target, model_points):
        pr, pt, pc, emb = self.estimator.forward_features(out_img, points, choose, obj)
        _, dis, new_points, new_target = self.criterion(pr, pt, pc, target, model_points, obj, points, self.w,
                                                        self.refine_start)
        if dis.dim() == 0:
            dis = dis.reshape(1)
        pose = self.criterion.last_pose.clone()          # (F,12) camera-frame pose after the best vote
        for it in range(self.refine_iters):
            # every iteration sees the same emb: from the second one on its branch is reused from the workspace
            rr, rt = self.refiner(new_points, emb, obj) if it == 0 else self.refiner.forward_next_iteration(new_points, emb, obj)
            dis, new_points, new_target = self.criterion_refine(rr, rt, new_target, model_points, obj, new_points)
            ops.pose_compose_(pose, self.criterion_refine.last_pose)    # stays on the device (kernel (5))
        return dis, pose

    def _param_signature(self):
        return tuple((p.data_ptr(), p._version) for m in (self.estimator, self.refiner) for p in m.hot_path_parameters())

    def _run_chunk(self, *args):
        """`_chunk`, through a captured CUDA graph when enabled.  The returned tensors of a graphed chunk live in the
        graph's memory pool and are overwritten by its next replay: callers copy them out (infer / infer_host do)."""
        if not self.cuda_graph:
            return self._chunk(*args)
        key = tuple((t.data_ptr(), tuple(t.shape), t.dtype) for t in args) + (self.refine_iters, self.w, self.refine_start,
                                                                              self._param_signature())
        g = self._graphs.get(key)
        if g is None:
            if len(self._graphs) >= self.max_graphs:
                self._graphs.pop(next(iter(self._graphs)))          # oldest capture
            cur = torch.cuda.current_stream(self.device)
            side = torch.cuda.Stream(device=self.device)
            side.wait_stream(cur)
            with torch.cuda.stream(side):                            # warm-up outside the capture: lazily created state
                self._chunk(*args)                                   # (packed weights, workspaces, LUTs, function attributes)
            cur.wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                out = self._chunk(*args)
            # keep the inputs and the workspaces the captured kernels point at alive with the graph
            g = (graph, out, args, (self.estimator._state.ws, self.refiner._state.ws))
            self._graphs[key] = g
        g[0].replay()
        return g[1]

    @torch.no_grad()
    def infer(self, out_img, points, choose, obj, target, model_points, return_pose=False):
        """Device-resident inputs (F, ...).  Returns the final ADD / ADD-S distance per frame (F,) and, with
        return_pose, the refined camera-frame poses (F,12) = R row-major | t."""
        F = out_img.shape[0]
        out = torch.empty(F, dtype=torch.float32, device=out_img.device)
        poses = torch.empty((F, 12), dtype=torch.float32, device=out_img.device)
        for a in range(0, F, self.chunk_frames):
            b = min(F, a + self.chunk_frames)
            out[a:b], poses[a:b] = self._run_chunk(out_img[a:b], points[a:b], choose[a:b], obj[a:b], target[a:b],
                                                   model_points[a:b])
        return (out, poses) if return_pose else out

    @torch.no_grad()
    def infer_host(self, host):
        """`host`: dict of PINNED CPU tensors with the INPUT_KEYS.  Chunks are copied on a side stream
        into two alternating device staging sets while the previous chunk computes; the per-frame
        result comes back to (pinned) host memory.  Returns a CPU tensor (F,)."""
        F = host["out_img"].shape[0]
        cf = min(self.host_chunk_frames, F)
        dev = self.device
        if self._staging is None or self._staging[0]["out_img"].shape[0] != cf or \
                any(self._staging[0][k].shape[1:] != host[k].shape[1:] for k in INPUT_KEYS):
            self._staging = [{k: torch.empty((cf,) + tuple(host[k].shape[1:]), dtype=host[k].dtype, device=dev)
                              for k in INPUT_KEYS} for _ in range(2)]
            self._copy_stream = torch.cuda.Stream(device=dev)
            self._copy_extra = [torch.cuda.Stream(device=dev) for _ in range(max(0, self.copy_streams - 1))]
            self._ready = [torch.cuda.Event() for _ in range(2)]
            self._free = [torch.cuda.Event() for _ in range(2)]
        res_dev = torch.empty(F, dtype=torch.float32, device=dev)
        pose_dev = torch.empty((F, 12), dtype=torch.float32, device=dev)
        res_host = torch.empty(F, dtype=torch.float32, pin_memory=True)
        pose_host = torch.empty((F, 12), dtype=torch.float32, pin_memory=True)
        main = torch.cuda.current_stream(dev)
        chunks = [(a, min(F, a + cf)) for a in range(0, F, cf)]

        def issue_copy(i):
            a, b = chunks[i]
            s = self._staging[i & 1]
            with torch.cuda.stream(self._copy_stream):
                if i >= 2:
                    self._copy_stream.wait_event(self._free[i & 1])   # compute of chunk i-2 has finished
                n = b - a
                parts = 1 + len(self._copy_extra)
                cut = [n * j // parts for j in range(parts + 1)]
                for k in INPUT_KEYS:                 # the large tensor (out_img) is split by frames over the copy streams
                    hi = cut[1] if (k == "out_img" and parts > 1) else n
                    s[k][:hi].copy_(host[k][a:a + hi], non_blocking=True)
                for j, xs in enumerate(self._copy_extra, start=1):
                    xs.wait_stream(self._copy_stream)          # ordering against the staging set's previous use
                    with torch.cuda.stream(xs):
                        s["out_img"][cut[j]:cut[j + 1]].copy_(host["out_img"][a + cut[j]:a + cut[j + 1]], non_blocking=True)
                    self._copy_stream.wait_stream(xs)
                self._ready[i & 1].record(self._copy_stream)

        issue_copy(0)
        for i, (a, b) in enumerate(chunks):
            if i + 1 < len(chunks):
                issue_copy(i + 1)
            main.wait_event(self._ready[i & 1])
            s = self._staging[i & 1]
            n = b - a
            res_dev[a:b], pose_dev[a:b] = self._run_chunk(*(s[k][:n] for k in INPUT_KEYS))
            self._free[i & 1].record(main)
        res_host.copy_(res_dev, non_blocking=True)
        pose_host.copy_(pose_dev, non_blocking=True)
        main.synchronize()
        self.last_poses = pose_host          # (F,12) refined poses of this call
        return res_host

    @staticmethod
    def bytes_per_frame(host):
        return sum(int(host[k][0].numel()) * host[k].element_size() for k in INPUT_KEYS)


def host_staging_buffer(shape, dtype, write_combined=False):
    """A page-locked host tensor for `infer_host`.  write_combined=True allocates it with cudaHostAllocWriteCombined:
    memory the CPU only WRITES (a staging buffer filled by the loader) and the GPU reads over PCIe without cache
    snooping - worth trying where several ranks stream from the same host (e2e at N = 8 is host-bound).
    CPU reads of such memory are slow: do not use it for results.  The storage is freed when the tensor is."""
    if not write_combined:
        return torch.empty(shape, dtype=dtype, pin_memory=True)
    import ctypes
    rt = ctypes.CDLL("libcudart.so.12")          # the runtime torch itself has loaded
    rt.cudaHostAlloc.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t, ctypes.c_uint]
    rt.cudaFreeHost.argtypes = [ctypes.c_void_p]
    nbytes = int(torch.empty((), dtype=dtype).element_size())
    for d in shape:
        nbytes *= int(d)
    ptr = ctypes.c_void_p()
    rc = rt.cudaHostAlloc(ctypes.byref(ptr), max(nbytes, 1), 0x04)          # cudaHostAllocWriteCombined
    if rc != 0:
        raise RuntimeError(f"cudaHostAlloc(write-combined, {nbytes} bytes) failed with cudaError {rc}")

    class _Owner:                                  # keeps the allocation alive as long as the tensor's buffer is
        def __init__(self, p):
            self.p = p
            self.buf = (ctypes.c_byte * max(nbytes, 1)).from_address(p.value)

        def __del__(self):
            rt.cudaFreeHost(self.p)
    owner = _Owner(ptr)
    t = torch.frombuffer(owner.buf, dtype=dtype, count=nbytes // torch.empty((), dtype=dtype).element_size()).reshape(shape)
    t._plr2_owner = owner
    return t
