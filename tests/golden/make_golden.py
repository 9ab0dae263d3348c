"""Generate tests/golden/*.npz by running the REFERENCE's own modules (imported from /root/reference;
nothing is copied) on seeded synthetic inputs.  Run where /root/reference exists:

    python tests/golden/make_golden.py

Shims needed to run the reference on torch 2.x / CPU, all kept out of the
reference tree:
  * sys.modules['lib.knn.knn_pytorch'] is pre-seeded with a module whose knn() calls the
    reference's OWN knn_cpu, compiled unmodified into oracle/_ref/libknn_cpu_ref.so;
  * lib.loss / lib.loss_refiner's KNearestNeighbor is rebound to a CPU shim that accepts `k=` as a
    keyword (lib/loss_refiner.py:43) and skips the unconditional .cuda() (lib/knn/__init__.py:14-17);
  * PoseNet.cnn (the out-of-scope PSPNet backbone) is replaced by nn.Identity so that the `img`
    argument carries the precomputed backbone output `out_img`.
Fixtures store only seeds, small inputs and the reference outputs; weights/frames are regenerated
from plr2_b200.synth by seed.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("PLR2_REFERENCE", "/root/reference")

from oracle import knn_oracle as KO  # noqa: E402
from plr2_b200 import synth  # noqa: E402


def import_reference():
    stub = types.ModuleType("lib.knn.knn_pytorch")

    def knn(ref, query, idx):
        for b in range(ref.shape[0]):
            out = KO.knn_reference_cpu(ref[b].detach().numpy(), query[b].detach().numpy(), idx.shape[1])
            idx[b].copy_(torch.from_numpy(out))
        return 1

    stub.knn = knn
    stub.knn_pytorch = stub
    sys.modules["lib.knn.knn_pytorch"] = stub
    sys.path.insert(0, REF)
    import lib.loss as rloss
    import lib.loss_refiner as rloss_ref
    import lib.network as rnet

    class KNNShim:
        @staticmethod
        def apply(ref, query, k=1):
            inds = torch.empty(query.shape[0], k, query.shape[2], dtype=torch.long)
            knn(ref.float().contiguous(), query.float().contiguous(), inds)
            return inds

    rloss.KNearestNeighbor = KNNShim
    rloss_ref.KNearestNeighbor = KNNShim
    return rnet, rloss, rloss_ref


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def load_params(module, params):
    sd = module.state_dict()
    for k, v in params.items():
        assert k in sd and tuple(sd[k].shape) == v.shape, k
    module.load_state_dict({k: t(v) for k, v in params.items()}, strict=False)


def sub(a, step=97):
    """strided subsample of a flattened array (keeps fixtures small)"""
    return np.ascontiguousarray(a.reshape(-1)[::step])


def case_posenet(rnet, rloss, name, shape, seed, obj, w=0.015, refine=False):
    torch.manual_seed(0)
    net = rnet.PoseNet(num_points=shape.num_points, num_obj=shape.num_obj)
    net.cnn = torch.nn.Identity()
    params = synth.make_params("posenet", shape.num_obj, seed)
    load_params(net, params)
    fr = synth.make_frame(shape, seed + 1, obj=obj)
    crit = rloss.Loss(shape.num_mesh, list(shape.sym_list))
    img = t(fr["out_img"])[None]
    points = t(fr["points"])[None]
    choose = t(fr["choose"])[None, None]
    idx = t(fr["idx"]).reshape(1, 1)
    pr, pt, pc, emb = net(img, points, choose, idx)
    pr.retain_grad(); pt.retain_grad(); pc.retain_grad()
    loss, dis, new_points, new_target = crit(pr, pt, pc, t(fr["target"])[None], t(fr["model_points"])[None],
                                             idx, points, w, refine)
    loss.backward()
    out = {
        "seed": seed, "obj": obj, "w": w, "refine": int(refine),
        "shape": np.array([shape.num_points, shape.num_obj, shape.num_mesh, shape.height, shape.width]),
        "sym_list": np.array(shape.sym_list),
        "pred_r": pr.detach().numpy(), "pred_t": pt.detach().numpy(), "pred_c": pc.detach().numpy(),
        "emb": emb.numpy(), "loss": loss.detach().numpy(), "dis": dis.detach().numpy(),
        "new_points": new_points.numpy(), "new_target": new_target.numpy(),
        "g_pred_r": pr.grad.numpy(), "g_pred_t": pt.grad.numpy(), "g_pred_c": pc.grad.numpy(),
    }
    for k, p in net.named_parameters():
        if p.grad is not None:
            out["gnorm." + k] = np.array(p.grad.norm().item(), dtype=np.float64)
            out["gsub." + k] = sub(p.grad.numpy())
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "loss", float(loss), "dis", float(dis))


def case_loss_only(rloss, name, shape, seed, obj, w=0.015, refine=False):
    fr = synth.make_frame(shape, seed, obj=obj, with_image=False)
    pr, pt, pc = [t(a)[None].requires_grad_(True) for a in synth.make_predictions(fr, seed + 7)]
    crit = rloss.Loss(shape.num_mesh, list(shape.sym_list))
    idx = t(fr["idx"]).reshape(1, 1)
    loss, dis, new_points, new_target = crit(pr, pt, pc, t(fr["target"])[None], t(fr["model_points"])[None],
                                             idx, t(fr["points"])[None], w, refine)
    (loss + 0.5 * dis).backward()   # exercises both differentiable outputs
    np.savez_compressed(os.path.join(HERE, name + ".npz"), seed=seed, obj=obj, w=w, refine=int(refine),
                        shape=np.array([shape.num_points, shape.num_obj, shape.num_mesh, shape.height, shape.width]),
                        sym_list=np.array(shape.sym_list), loss=loss.detach().numpy(), dis=dis.detach().numpy(),
                        new_points=new_points.numpy(), new_target=new_target.numpy(),
                        g_pred_r=pr.grad.numpy(), g_pred_t=pt.grad.numpy(), g_pred_c=pc.grad.numpy())
    print(name, "loss", float(loss), "dis", float(dis))


def case_refiner(rnet, rloss_ref, name, shape, seed, obj):
    torch.manual_seed(0)
    net = rnet.PoseRefineNet(num_points=shape.num_points, num_obj=shape.num_obj)
    params = synth.make_params("refiner", shape.num_obj, seed)
    load_params(net, params)
    fr = synth.make_frame(shape, seed + 1, obj=obj)
    rng = np.random.Generator(np.random.PCG64(seed + 2))
    # a plausible hand-off from the estimator stage: re-centred points / target in the object frame
    rot, _ = synth.random_rotation(rng)
    tt = fr["gt_t"] + rng.normal(0, 0.01, 3).astype(np.float32)
    new_points = ((fr["points"] - tt) @ rot).astype(np.float32)
    new_target = ((fr["target"] - tt) @ rot).astype(np.float32)
    emb = fr["out_img"].reshape(32, -1)[:, fr["choose"]]
    crit = rloss_ref.Loss_refine(shape.num_mesh, list(shape.sym_list))
    idx = t(fr["idx"]).reshape(1, 1)
    out = {"seed": seed, "obj": obj, "sym_list": np.array(shape.sym_list),
           "shape": np.array([shape.num_points, shape.num_obj, shape.num_mesh, shape.height, shape.width]),
           "in_points": new_points, "in_target": new_target}
    pts, tgt = t(new_points)[None], t(new_target)[None]
    for ite in range(2):   # tools/train.py:141-145
        pr, pt = net(pts, t(emb)[None], idx)
        pr.retain_grad(); pt.retain_grad()
        dis, pts, tgt = crit(pr, pt, tgt, t(fr["model_points"])[None], idx, pts)
        dis.backward()
        out[f"it{ite}.pred_r"] = pr.detach().numpy(); out[f"it{ite}.pred_t"] = pt.detach().numpy()
        out[f"it{ite}.g_pred_r"] = pr.grad.numpy(); out[f"it{ite}.g_pred_t"] = pt.grad.numpy()
        out[f"it{ite}.dis"] = dis.detach().numpy()
        out[f"it{ite}.new_points"] = pts.numpy(); out[f"it{ite}.new_target"] = tgt.numpy()
    for k, p in net.named_parameters():   # accumulated over both iterations, like tools/train.py
        out["gnorm." + k] = np.array(p.grad.norm().item(), dtype=np.float64)
        out["gsub." + k] = sub(p.grad.numpy())
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "dis", [float(out[f"it{i}.dis"][0]) for i in range(2)])


def case_knn(name, seed, d, n_ref, n_query, k):
    rng = np.random.Generator(np.random.PCG64(seed))
    ref = rng.standard_normal((d, n_ref)).astype(np.float32)
    query = rng.standard_normal((d, n_query)).astype(np.float32)
    if d == 3:  # exact ties: duplicate some reference points and snap some queries onto references
        ref[:, 5] = ref[:, 2]
        ref[:, 11] = ref[:, 2]
        query[:, :8] = ref[:, [2, 5, 11, 0, 1, 3, 4, 2]]
    idx = KO.knn_reference_cpu(ref, query, k)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), seed=seed, ref=ref, query=query, k=k, idx=idx)
    print(name, idx.shape)


def main():
    rnet, rloss, rloss_ref = import_reference()
    small_lm = synth.Shape(96, 13, 60, 16, 16, tuple(synth.LINEMOD_SYM))
    small_ycb = synth.Shape(128, 21, 70, 20, 20, tuple(synth.YCB_SYM))
    case_posenet(rnet, rloss, "posenet_linemod_nonsym", small_lm, 100, obj=3)
    case_posenet(rnet, rloss, "posenet_ycb_sym", small_ycb, 200, obj=12)
    case_posenet(rnet, rloss, "posenet_ycb_refine", small_ycb, 300, obj=12, refine=True)
    case_loss_only(rloss, "loss_ycb_sym", synth.Shape(200, 21, 120, 20, 20, tuple(synth.YCB_SYM)), 400, obj=15)
    case_loss_only(rloss, "loss_ycb_nonsym", synth.Shape(200, 21, 120, 20, 20, tuple(synth.YCB_SYM)), 410, obj=3)
    case_loss_only(rloss, "loss_linemod_sym", synth.Shape(150, 13, 90, 16, 16, tuple(synth.LINEMOD_SYM)), 420, obj=8)
    case_refiner(rnet, rloss_ref, "refiner_ycb_sym", small_ycb, 500, obj=18)
    case_refiner(rnet, rloss_ref, "refiner_ycb_nonsym", small_ycb, 510, obj=4)
    case_knn("knn_d3_k1", 600, 3, 300, 2000, 1)
    case_knn("knn_d3_k4", 601, 3, 64, 500, 4)
    case_knn("knn_d16_k2", 602, 16, 100, 300, 2)   # the (scaled-down) shape of the reference's own unit test


if __name__ == "__main__":
    main()
