"""GPU: plr2_refine_loss_fwd_static - the refiner loss with its scratch in a static device array inside the library
(no counterpart in the reference; added for callers that cannot allocate).
Called through the C-ABI and compared with the existing entry point plr2_refine_loss_fwd on the same inputs, on two
calls in a row (the shared array keeps the first call's contents: the second call - fewer frames, other classes, other
mesh size - must not see them).  Collected last on purpose."""
import numpy as np
import pytest
import torch

from plr2_b200 import _cabi, ops, synth
from tests.util import assert_close

pytestmark = pytest.mark.gpu

OUT = ("dis", "new_points", "new_target", "jac_r", "jac_t", "pose")


def _inputs(frames, n_mesh, n_points, seed):
    shp = synth.Shape(n_points, 21, n_mesh, 8, 8, tuple(synth.YCB_SYM))
    objs = [12, 3, 19, 0, 7, 20, 5, 16]
    base = [synth.make_frame(shp, seed + i, obj=o, with_image=False) for i, o in enumerate(objs)]
    pick = [(i * 3 + seed) % 8 for i in range(frames)]
    rng = np.random.default_rng(seed)
    t = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to("cuda", dt).contiguous()
    return {
        "pred_r": t(np.stack([base[i]["gt_q"] + 0.05 * rng.standard_normal(4) for i in pick])),
        "pred_t": t(np.stack([base[i]["gt_t"] + 0.005 * rng.standard_normal(3) for i in pick])),
        "target": t(np.stack([base[i]["target"] for i in pick])),
        "model": t(np.stack([base[i]["model_points"] for i in pick])),
        "points": t(np.stack([base[i]["points"] for i in pick])),
        "obj": t(np.array([int(base[i]["idx"]) for i in pick]), torch.int64),
        "lut": ops.sym_lut(shp.sym_list, torch.device("cuda", torch.cuda.current_device())),
        "F": frames, "N": n_points, "M": n_mesh,
    }


def _outputs(x):
    F, N, M = x["F"], x["N"], x["M"]
    mk = lambda *s: torch.full(s, float("nan"), dtype=torch.float32, device="cuda")
    return {"dis": mk(F), "new_points": mk(F, N, 3), "new_target": mk(F, M, 3), "jac_r": mk(F, 4), "jac_t": mk(F, 3),
            "pose": mk(F, 12)}


def _call(x, mode, static):
    L = _cabi.lib()
    o = _outputs(x)
    p = _cabi.ptr
    st = _cabi.stream_ptr()
    common = (p(x["pred_r"]), p(x["pred_t"]), p(x["target"]), p(x["model"]), p(x["points"]), p(x["obj"]), p(x["lut"]),
              int(x["lut"].numel()), x["F"], x["N"], x["M"], int(mode), p(o["dis"]), p(o["new_points"]), p(o["new_target"]),
              p(o["jac_r"]), p(o["jac_t"]), p(o["pose"]))
    if static:
        rc = L.plr2_refine_loss_fwd_static(*common, st)
        _cabi.check(rc, "plr2_refine_loss_fwd_static")
    else:
        nws = L.plr2_refine_loss_workspace_bytes(x["F"], x["M"])
        ws = torch.empty(nws, dtype=torch.uint8, device="cuda")
        rc = L.plr2_refine_loss_fwd(*common, p(ws), nws, st)
        _cabi.check(rc, "plr2_refine_loss_fwd")
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in o.items()}


@pytest.mark.parametrize("mode", [ops.KNN_EXACT, ops.KNN_FAST])
def test_static_scratch_entry_matches_the_workspace_entry_on_two_calls_in_a_row(mode):
    first = _inputs(200, 2300, 64, 11)        # QPT = 4 instantiation, 24 slices per frame, 75 symmetric frames
    second = _inputs(37, 500, 40, 12)         # fewer frames and slices than the call before: stale partials / counters beyond them
    for x in (first, second):
        got = _call(x, mode, static=True)
        want = _call(x, mode, static=False)
        for k in OUT:
            assert np.isfinite(got[k]).all(), k
            assert_close(got[k], want[k], k, frac=1.0)      # tolerance of tests/util.py (rtol 1e-3 + 1e-4 max|b|)
    # and once more with the first inputs: same answer as before (the array is re-armed by every call)
    again = _call(first, mode, static=True)
    want = _call(first, mode, static=False)
    for k in OUT:
        assert_close(again[k], want[k], k + " (third call)", frac=1.0)


def test_static_scratch_entry_limits_and_size():
    L = _cabi.lib()
    assert L.plr2_refine_loss_static_scratch_bytes() == L.plr2_refine_loss_workspace_bytes(256, 8192) == 427008
    x = _inputs(2, 100, 16, 5)
    o = _outputs(x)
    p = _cabi.ptr
    args = lambda F, M: (p(x["pred_r"]), p(x["pred_t"]), p(x["target"]), p(x["model"]), p(x["points"]), p(x["obj"]), p(x["lut"]),
                         int(x["lut"].numel()), F, x["N"], M, 0, p(o["dis"]), p(o["new_points"]), p(o["new_target"]),
                         p(o["jac_r"]), p(o["jac_t"]), p(o["pose"]), _cabi.stream_ptr())
    for F, M in ((257, 100), (2, 8193)):      # beyond the stated maximum: rejected before anything is launched
        assert L.plr2_refine_loss_fwd_static(*args(F, M)) == -3
        msg = L.plr2_last_error_string()
        assert b"beyond the static scratch" in msg and b"plr2_refine_loss_fwd" in msg
    assert L.plr2_refine_loss_fwd_static(*args(0, 100)) == 0          # empty batch, like the workspace entry
    torch.cuda.synchronize()
    assert bool(torch.isnan(o["dis"]).all())                          # nothing was written by the rejected / empty calls
