"""GPU: CUDA-graph replay of the launch chains (plr2_b200/pipeline.py, plr2_b200/train.py) gives exactly what eager
launches give - same kernels, same order - including when inputs, shapes or parameters change between calls."""
import numpy as np
import pytest
import torch

from plr2_b200 import ops, synth
from plr2_b200.pipeline import DenseFusionPipeline, INPUT_KEYS
from tests.util import tt

pytestmark = pytest.mark.gpu


def _work(shp, objs, seed):
    frs = [synth.make_frame(shp, seed + i, obj=o) for i, o in enumerate(objs)]
    st = lambda k: tt(np.stack([f[k] for f in frs]), "cuda")
    return {"out_img": st("out_img"), "points": st("points"), "choose": st("choose")[:, None], "obj": st("idx")[:, None],
            "target": st("target"), "model_points": st("model_points")}


def _pipe(shp, graph, chunk):
    pipe = DenseFusionPipeline(shp.num_points, shp.num_obj, shp.num_mesh, shp.sym_list, device="cuda:0",
                               precision=ops.PREC_BF16X2, chunk_frames=chunk, refine_iters=2, cuda_graph=graph, max_graphs=4)
    pipe.load_params(synth.make_params("posenet", shp.num_obj, 31), synth.make_params("refiner", shp.num_obj, 32))
    return pipe


def test_pipeline_graph_replay_equals_eager_launches():
    shp = synth.Shape(1000, 21, 600, 40, 40, tuple(synth.YCB_SYM))
    eager, graph = _pipe(shp, False, 4), _pipe(shp, True, 4)
    a = _work(shp, [12, 3, 19, 0, 7, 20], 700)          # 6 frames -> chunks of 4 + 2: two graphs
    b = _work(shp, [5, 15, 1, 18, 2, 9], 800)
    for w in (a, b, a):                                  # b has other addresses: new captures; a again: replays of the first
        de, pe = eager.infer(*(w[k] for k in INPUT_KEYS), return_pose=True)
        for _ in range(2):
            dg, pg = graph.infer(*(w[k] for k in INPUT_KEYS), return_pose=True)
            assert torch.equal(de, dg) and torch.equal(pe, pg)
    assert len(graph._graphs) == 4
    # in-place change of an input the graph reads: the replay must see the new data
    a["points"].add_(0.001)
    de = eager.infer(*(a[k] for k in INPUT_KEYS))
    dg = graph.infer(*(a[k] for k in INPUT_KEYS))
    assert torch.equal(de, dg)
    # parameter update (version bump): stale graphs must not be replayed
    for p in (eager, graph):
        with torch.no_grad():
            p.refiner.conv3_t.bias.add_(0.01)
    de = eager.infer(*(a[k] for k in INPUT_KEYS))
    dg = graph.infer(*(a[k] for k in INPUT_KEYS))
    assert torch.equal(de, dg)


def test_infer_host_with_graphs_equals_eager():
    shp = synth.Shape(1000, 21, 600, 40, 40, tuple(synth.YCB_SYM))
    w = _work(shp, [12, 3, 19, 0, 7, 20, 5, 16, 1, 2], 900)
    host = {k: v.cpu().pin_memory() for k, v in w.items()}
    outs = []
    for graph in (False, True):
        pipe = _pipe(shp, graph, 4)
        pipe.host_chunk_frames = 4                       # chunks of 4, 4, 2 through the two staging sets
        r1 = pipe.infer_host(host).clone()
        r2 = pipe.infer_host(host).clone()
        assert torch.equal(r1, r2)
        outs.append((r1, pipe.last_poses.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


def test_c5_train_step_graph_equals_eager():
    """The whole training step (forward, loss, backward, gradient bucket, fused Adam) replayed as ONE CUDA graph computes
    what the eager step computes: from equal parameters the first replay returns the eager step's losses bit for bit and
    its gradients to round-off (the few FP32-SIMT weight gradients sum split-K partials with atomics)."""
    from plr2_b200.train import C5TrainStep
    shp = synth.Shape(256, 21, 300, 40, 40, tuple(synth.YCB_SYM))
    mk = lambda: C5TrainStep("cuda:0", 1, 0, 4, precision=ops.PREC_BF16X2_ALL, cuda_graph=False, shape=shp, lr=1e-3)
    ts_e = mk()
    loss_e = ts_e.step().detach().clone()
    grads_e = ts_e.bucket.flat.clone()
    ts_g = mk()
    ts_g.opt = torch.optim.Adam(ts_g.params, lr=1e-3, fused=True, capturable=True)
    start = [p.detach().clone() for p in ts_g.params]
    assert ts_g.capture(warmup=3), ts_g.graph_error
    with torch.no_grad():                                # undo the warm-up / capture updates: same parameters as ts_e had
        for p, s0 in zip(ts_g.params, start):
            p.copy_(s0)
    loss_g = ts_g.step().detach().clone()
    grads_g = ts_g.bucket.flat.clone()
    torch.cuda.synchronize()
    assert ts_g.sym_frames == ts_e.sym_frames == 1
    assert torch.equal(loss_e, loss_g)
    scale = float(grads_e.abs().max())
    assert float((grads_e - grads_g).abs().max()) <= 1e-5 * scale
    before = [p.detach().clone() for p in ts_g.params]
    for _ in range(3):                                   # further replays keep training (parameters move, losses stay finite)
        l = ts_g.step()
    torch.cuda.synchronize()
    assert bool(torch.isfinite(l).all()) and any(not torch.equal(a, b) for a, b in zip(before, ts_g.params))


def test_infer_host_from_write_combined_staging_buffers():
    """pipeline.host_staging_buffer(write_combined=True): page-locked write-combined host memory is a valid source of
    the host path (recognised as pinned, same results as ordinary pinned buffers) and is released with the tensor."""
    from plr2_b200.pipeline import host_staging_buffer
    shp = synth.Shape(1000, 21, 600, 40, 40, tuple(synth.YCB_SYM))
    w = _work(shp, [12, 3, 19, 0, 7], 950)
    pinned = {k: v.cpu().pin_memory() for k, v in w.items()}
    wc = {}
    for k, v in w.items():
        wc[k] = host_staging_buffer(tuple(v.shape), v.dtype, write_combined=True)
        assert wc[k].is_pinned() and tuple(wc[k].shape) == tuple(v.shape) and wc[k].dtype == v.dtype
        wc[k].copy_(v.cpu())
    pipe = _pipe(shp, False, 4)
    pipe.host_chunk_frames = 2
    a = pipe.infer_host(pinned).clone()
    b = pipe.infer_host(wc).clone()
    assert torch.equal(a, b)
    del wc
