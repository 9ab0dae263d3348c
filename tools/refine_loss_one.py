"""One Loss_refine launch shape for ncu captures: refine_loss_one.py <frames> [mode]  (5 of 16 frames symmetric, M = 2300)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from plr2_b200 import ops, synth
from plr2_b200.lib.loss_refiner import Loss_refine
F = int(sys.argv[1]) if len(sys.argv) > 1 else 64
mode = int(sys.argv[2]) if len(sys.argv) > 2 else ops.KNN_FAST
shp = synth.YCB_REFINE
objs = [12, 3, 0, 19, 1, 2, 4, 15, 5, 6, 7, 18, 8, 9, 10, 20]
base = [synth.make_frame(shp, 100 + i, obj=o, with_image=False) for i, o in enumerate(objs)]
pick = [i % 16 for i in range(F)]
st = lambda k: torch.from_numpy(np.stack([base[i][k] for i in pick])).cuda()
rng = np.random.default_rng(0)
pr = torch.from_numpy(np.stack([base[i]["gt_q"] + 0.05 * rng.standard_normal(4) for i in pick]).astype(np.float32)).cuda()
pt = torch.from_numpy(np.stack([base[i]["gt_t"] + 0.005 * rng.standard_normal(3) for i in pick]).astype(np.float32)).cuda()
cr = Loss_refine(shp.num_mesh, list(shp.sym_list), mode)
tgt, mdl, ob, pts = st("target"), st("model_points"), st("idx")[:, None], st("points")
with torch.no_grad():
    for _ in range(4):
        cr(pr, pt, tgt, mdl, ob, pts)
torch.cuda.synchronize()
print("ok")
