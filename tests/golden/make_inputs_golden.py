"""Generates tests/golden/inputs_ycb.npz by running the REFERENCE's own loader - datasets/ycb/dataset.py:
PoseDataset.__init__ / __getitem__ (:19-232) and get_bbox (:251-289), imported from /root/reference - on a small
synthetic YCB-Video tree written to a temporary directory (PNG colour / depth / label, -meta.mat, models/*/points.xyz).
Run in the build container only (the GPU box has no /root/reference):  python tests/golden/make_inputs_golden.py

Pins row f-2 of SURVEY.md 8: object mask, crop box, `choose`, depth back-projection, model sampling hand-off and
target transform.  The raw frames are regenerated from seeds by the tests (tests/golden/inputs_synth.py); stored are the
reference's OUTPUTS (cloud, choose, target, model_points, idx), its crop box, and the two RNG-dependent internals
needed to replay it deterministically: which object it drew and which mask pixels its shuffled c_mask kept.
"""
import os
import random
import sys
import tempfile

import numpy as np
import scipy.io as scio
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import inputs_synth as S  # noqa: E402

REF = "/root/reference"


def write_tree(root):
    classes = [l.strip() for l in open(os.path.join(REF, "datasets/ycb/dataset_config/classes.txt")) if l.strip()]
    rng = np.random.Generator(np.random.PCG64(99))
    for ci, name in enumerate(classes):
        os.makedirs(os.path.join(root, "models", name), exist_ok=True)
        pts = (rng.random((2800 + 7 * ci, 3)) - 0.5) * np.array([0.12, 0.09, 0.2])
        with open(os.path.join(root, "models", name, "points.xyz"), "w") as f:
            for p in pts:
                f.write("%.6f %.6f %.6f\n" % tuple(p))
    for name, seed, blobs in S.FRAMES:
        fr = S.raw_frame(seed, blobs)
        base = os.path.join(root, name)
        os.makedirs(os.path.dirname(base), exist_ok=True)
        Image.fromarray(fr["color"]).save(base + "-color.png")
        Image.fromarray(fr["depth"]).save(base + "-depth.png")
        Image.fromarray(fr["label"]).save(base + "-label.png")
        scio.savemat(base + "-meta.mat", {"poses": fr["poses"], "cls_indexes": fr["cls_indexes"],
                                          "factor_depth": fr["factor_depth"]})


def main():
    root = tempfile.mkdtemp(prefix="plr2_ycb_")
    write_tree(root)
    os.chdir(REF)                                   # the loader opens 'datasets/ycb/dataset_config/...' relative to cwd
    sys.path.insert(0, REF)
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_ycb_dataset", os.path.join(REF, "datasets/ycb/dataset.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    out = {}
    for refine, num_pt in ((False, 1000), (True, 500)):
        ds = mod.PoseDataset("test", num_pt, False, root, 0.0, refine)
        for k, (name, seed, blobs) in enumerate(S.FRAMES):
            index = ds.list.index(name)
            for rep in range(2):                    # two RNG states per frame: different objects / subsets
                np.random.seed(1000 * seed + rep)
                random.seed(2000 * seed + rep)
                cloud, choose, img, target, model_points, idx = ds[index]
                fr = S.raw_frame(seed, blobs)
                obj_id = int(idx[0]) + 1
                mask_label = fr["label"] == obj_id
                bbox = mod.get_bbox(mask_label)     # the reference's own crop box for the object it drew
                key = f"f{k}_r{int(refine)}_{rep}"
                out[key + "_cloud"] = cloud.numpy()
                out[key + "_choose"] = choose.numpy().astype(np.int64)
                out[key + "_target"] = target.numpy()
                out[key + "_model_points"] = model_points.numpy()
                out[key + "_idx"] = np.int64(int(idx[0]))
                out[key + "_bbox"] = np.array(bbox, dtype=np.int64)
                out[key + "_num_pt"] = np.int64(num_pt)
                which = [b[0] for b in blobs].index(obj_id)
                out[key + "_pose_col"] = np.int64(which)
                assert tuple(img.shape[1:]) == (bbox[1] - bbox[0], bbox[3] - bbox[2])
    path = os.path.join(HERE, "inputs_ycb.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
