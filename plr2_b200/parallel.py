"""Host-side multi-GPU logic: one process per GPU, frames sharded, weights replicated.

Frames are independent, so inference needs no data-path collective — each rank runs its shard and
the per-frame results are all-gathered once at the end.  Training adds exactly one collective per
optimiser step: the gradient all-reduce (NCCL over NVLink on the GPU box; the same code runs on the
gloo backend in the CPU tests).  The reference has no distributed code at all; the
accumulation semantics follow its `accumulate_grad_batches: 8` (yaml/exp/exp_load_and_fit.yml:13):
Lightning AVERAGES the accumulated frames, tools/train.py:152-155 SUMS them — `average` selects which.
"""
import torch
import torch.distributed as dist


def shard_range(n_frames, world_size, rank):
    """Contiguous, balanced shard [a, b) of `n_frames` for `rank` (the first n % world ranks get one extra)."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank {rank} / world {world_size}")
    base, extra = divmod(int(n_frames), world_size)
    a = rank * base + min(rank, extra)
    return a, a + base + (1 if rank < extra else 0)


def shard_frames(batch, world_size, rank):
    """Slice every (F, ...) tensor of a dict to this rank's frames."""
    n = next(iter(batch.values())).shape[0]
    a, b = shard_range(n, world_size, rank)
    return {k: v[a:b] for k, v in batch.items()}


def frame_cost(obj, sym_list, n_points, n_mesh):
    """Relative cost of one training frame in pair evaluations: ADD-S frames search the nearest of `n_mesh` targets
    for each of the n_points * n_mesh transformed model points (lib/loss.py:39-46), ADD frames evaluate n_points *
    n_mesh matched pairs (lib/loss.py:48); the network part is the same for every frame (~4 GFLOP on tensor cores,
    a small constant next to either)."""
    pairs = n_points * n_mesh
    return float(pairs * n_mesh) if int(obj) in sym_list else float(pairs)


def deal_by_cost(costs, world_size):
    """Cost-aware assignment of the frames of a global batch to ranks: frames sorted by decreasing cost (ties by
    index) are dealt to the currently lightest rank that still has room, every rank ending with the same number of
    frames (+-1).  Frames are independent, so any assignment is a valid data-parallel step; this one keeps the
    ranks' step times level (the all-reduce waits for the slowest rank).  Deterministic: every rank computes the
    same table.  Returns a list of `world_size` index lists (each ascending)."""
    n = len(costs)
    if world_size <= 0:
        raise ValueError("world_size must be positive")
    cap = [(n // world_size) + (1 if r < n % world_size else 0) for r in range(world_size)]
    order = sorted(range(n), key=lambda i: (-float(costs[i]), i))
    load = [0.0] * world_size
    out = [[] for _ in range(world_size)]
    for i in order:
        r = min((r for r in range(world_size) if len(out[r]) < cap[r]), key=lambda r: (load[r], r))
        out[r].append(i)
        load[r] += float(costs[i])
    return [sorted(ix) for ix in out]


def gather_results(local, n_frames, group=None):
    """All-gather per-frame results (f_local, ...) of uneven shards into (n_frames, ...) on every rank."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    sizes = [shard_range(n_frames, world, r) for r in range(world)]
    longest = max(b - a for a, b in sizes)
    pad = torch.zeros((longest,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    return torch.cat([p[: b - a] for p, (a, b) in zip(parts, sizes)], 0)


class GradBucket:
    """One flat fp32 buffer aliasing the gradients of a parameter list, so that the whole network's
    gradient is ONE all-reduce (PoseNet hot path: 3.99 M floats = 16 MB; refiner 1.94 M = 7.8 MB) —
    sized for launch latency, not link count (NVSwitch gives every pair full bandwidth)."""

    def __init__(self, params):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("GradBucket: no trainable parameters")
        dev, dt = self.params[0].device, self.params[0].dtype
        self.numel = sum(p.numel() for p in self.params)
        self.flat = torch.zeros(self.numel, dtype=dt, device=dev)
        o = 0
        for p in self.params:           # p.grad becomes a view into the flat buffer
            n = p.numel()
            view = self.flat[o:o + n].view_as(p)
            if p.grad is not None:
                view.copy_(p.grad)
            p.grad = view
            o += n

    def zero(self):
        self.flat.zero_()

    def all_reduce(self, average=True, group=None, async_op=False):
        """Sum (or average) the bucket over the ranks.  No-op without an initialised process group."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
        if average:
            if async_op:
                work.wait()
                work = None
            self.flat.div_(dist.get_world_size(group))
        return work


class InOpGradSync:
    """Gradient all-reduce performed INSIDE PoseNet's backward op, overlapped with it: attach with
    ``attach(net)``; the op then returns gradients that are already summed (averaged) over the ranks, so the caller
    must not all-reduce them again.  The packed gradient is reduced in three pieces: the head weights
    (``plr2_net_grad_regions``: 83 % of the bytes) on a side stream as soon as the backward's stage event fires,
    the trunk weights and the biases on the main stream when the backward has finished."""

    def __init__(self, average=True, group=None):
        self.average, self.group = average, group
        self._stream = None
        self._event = None

    def attach(self, net):
        net._state.grad_sync = self
        return self

    def active(self):
        return dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1

    def arm(self, dev):
        """Returns the raw cudaEvent_t the next backward records when the early region is complete."""
        if self._stream is None:
            self._stream = torch.cuda.Stream(device=dev)
            self._event = torch.cuda.Event()
            self._event.record(torch.cuda.current_stream(dev))          # forces creation of the CUDA event
        return self._event.cuda_event

    def reduce(self, gp, early_begin, early_end):
        dev = gp.device
        main = torch.cuda.current_stream(dev)
        early = gp[early_begin:early_end]
        self._stream.wait_event(self._event)                            # recorded by plr2_posenet_backward
        gp.record_stream(self._stream)
        with torch.cuda.stream(self._stream):
            dist.all_reduce(early, op=dist.ReduceOp.SUM, group=self.group)
        if early_begin > 0:
            dist.all_reduce(gp[:early_begin], op=dist.ReduceOp.SUM, group=self.group)
        if early_end < gp.numel():
            dist.all_reduce(gp[early_end:], op=dist.ReduceOp.SUM, group=self.group)
        main.wait_stream(self._stream)
        if self.average:
            gp.div_(dist.get_world_size(self.group))


def data_parallel_step(frames, step_fn, bucket, optimizer, average=True, group=None):
    """One data-parallel optimiser step: `step_fn(frame)` must run forward + backward of ONE frame
    (gradients accumulate into the bucket, like `accumulate_grad_batches`), then the single gradient
    all-reduce, then the optimiser.  Returns the mean of the per-frame losses seen by this rank."""
    bucket.zero()
    total, count = 0.0, 0
    for fr in frames:
        loss = step_fn(fr)
        total += float(loss)
        count += 1
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    if average and count:
        bucket.flat.div_(count)
    bucket.all_reduce(average=average, group=group)
    optimizer.step()
    return total / max(count, 1), world


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def bind_to_gpu_numa(device_index):
    """Restrict this process to the CPUs local to GPU `device_index` (sysfs `local_cpulist` of its PCI
    function), so that pinned staging buffers allocated afterwards land on the NUMA node the GPU's PCIe
    root hangs off.  With one rank per GPU, eight ranks each streaming ~50 GB/s from host memory otherwise
    share one node's DRAM and the inter-socket link.  Returns the CPU set applied, or None when the
    topology is not visible (containers without sysfs PCI, non-Linux): binding is an optimisation only."""
    import os
    try:
        p = torch.cuda.get_device_properties(device_index)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        with open("/sys/bus/pci/devices/%s/local_cpulist" % bdf) as f:
            cpus = _parse_cpulist(f.read())
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if not cpus or cpus == allowed:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except (OSError, AttributeError, ValueError, RuntimeError):
        return None
