"""GPU: evaluation metrics after the path (ADD / ADD-S between two poses, accuracy-threshold AUC) through the
C-ABI against the reference's own outputs (golden) and the float64 oracle."""
import numpy as np
import pytest
import torch

from oracle import metrics_oracle as MO
from plr2_b200 import _cabi, metrics
from tests.util import load_golden

pytestmark = pytest.mark.gpu

RTOL = 1e-3          # north_star tolerance for floating-point results; observed error is ~1e-6 (fp32 vs float64)


def _cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()


@pytest.mark.parametrize("name", ["small", "ycb", "far"])
def test_add_adds_match_reference_golden(name):
    g = load_golden("metrics_add_adds")
    add, adds = metrics.pose_metrics(_cuda(g[name + "_model"]), _cuda(g[name + "_pose1"]), _cuda(g[name + "_pose2"]))
    np.testing.assert_allclose(add.cpu().numpy(), g[name + "_add"], rtol=RTOL, atol=1e-7)
    np.testing.assert_allclose(adds.cpu().numpy(), g[name + "_adds"], rtol=RTOL, atol=1e-7)
    # the single-metric entry points (one output NULL) give the same numbers
    assert torch.equal(metrics.ADD(_cuda(g[name + "_model"]), _cuda(g[name + "_pose1"]), _cuda(g[name + "_pose2"])), add)
    assert torch.equal(metrics.ADDS(_cuda(g[name + "_model"]), _cuda(g[name + "_pose1"]), _cuda(g[name + "_pose2"])), adds)


@pytest.mark.parametrize("m", [1, 15, 16, 17, 500, 2300, 2600])
def test_shared_model_ragged_sizes_vs_oracle(m):
    rng = np.random.default_rng(m)
    frames = 5
    model = (rng.random((m, 3)) - 0.5) * 0.3
    def pose():
        q = rng.normal(size=4); q /= np.linalg.norm(q)
        w, x, y, z = q
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                      [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                      [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
        return np.concatenate([R.reshape(-1), rng.normal(size=3) * 0.05])
    p1 = np.stack([pose() for _ in range(frames)]); p2 = np.stack([pose() for _ in range(frames)])
    add, adds = metrics.pose_metrics(_cuda(model), _cuda(p1), _cuda(p2))          # (F,12) layout, shared model
    def T(p):
        M = np.eye(4); M[:3, :3] = p[:9].reshape(3, 3); M[:3, 3] = p[9:]; return M
    ref_add = [MO.add(model, T(p1[f]), T(p2[f])) for f in range(frames)]
    ref_adds = [MO.adds(model, T(p1[f]), T(p2[f])) for f in range(frames)]
    np.testing.assert_allclose(add.cpu().numpy(), ref_add, rtol=RTOL, atol=1e-7)
    np.testing.assert_allclose(adds.cpu().numpy(), ref_adds, rtol=RTOL, atol=1e-7)


def test_known_answers_and_identity():
    sym = np.array([[-1, 0, 0], [1, 0, 0], [0, -1, 0], [0, 1, 0]], dtype=np.float64)
    def rz(a):
        M = np.eye(4); c, s = np.cos(a), np.sin(a); M[:2, :2] = [[c, -s], [s, c]]; return M
    add, adds = metrics.pose_metrics(_cuda(sym), _cuda(rz(np.pi / 2)[None]), _cuda(rz(-np.pi / 2)[None]))
    assert abs(float(adds[0])) < 1e-6 and abs(float(add[0]) - 2.0) < 1e-6       # src/estimation/main.py:121-131
    p = _cuda(rz(0.3)[None])
    add, adds = metrics.pose_metrics(_cuda(sym), p, p)
    assert float(add[0]) == 0.0 and float(adds[0]) == 0.0                          # main.py:117-119


@pytest.mark.parametrize("case", ["random", "ties", "misses", "all_miss", "one", "perfect", "large"])
def test_auc_matches_oracle(case):
    rng = np.random.default_rng(11)
    d = {"random": rng.random(777) * 0.15,
         "ties": np.round(rng.random(400) * 0.12, 2),
         "misses": np.concatenate([rng.random(50) * 0.1, [np.inf, np.nan, 0.5, 0.1, 0.0, 0.0]]),
         "all_miss": np.array([0.2, np.inf, np.nan]),
         "one": np.array([0.03]),
         "perfect": np.zeros(9),
         "large": rng.random(20000) * 0.11}[case].astype(np.float32)
    out = metrics.auc(torch.from_numpy(d).cuda()).cpu().numpy()
    ref = MO.auc(d.astype(np.float64))
    assert out[0] == pytest.approx(ref[0], rel=1e-12, abs=1e-15)
    assert out[1] == pytest.approx(ref[1], rel=1e-12, abs=1e-15)


def test_auc_empty_and_other_threshold():
    out = metrics.auc(torch.empty(0, device="cuda")).cpu().numpy()
    assert out[0] == 0.0 and out[1] == 0.0
    d = np.linspace(0.0, 0.3, 301).astype(np.float32)
    out = metrics.auc(torch.from_numpy(d).cuda(), max_distance=0.2, below=0.05).cpu().numpy()
    ref = MO.auc(d.astype(np.float64), 0.2, 0.05)
    assert out[0] == pytest.approx(ref[0], rel=1e-12) and out[1] == pytest.approx(ref[1], rel=1e-12)


def test_pipeline_poses_feed_the_metrics():
    """ADD of the pipeline's composed pose against the ground-truth pose equals, for a non-symmetric frame,
    the mean distance between the transformed model points - i.e. the two layouts agree."""
    rng = np.random.default_rng(5)
    model = _cuda((rng.random((500, 3)) - 0.5) * 0.2)
    I = torch.eye(3, device="cuda").reshape(1, 9)
    pa = torch.cat([I, torch.tensor([[0.01, 0.0, 0.0]], device="cuda")], dim=1)
    pb = torch.cat([I, torch.zeros(1, 3, device="cuda")], dim=1)
    add, adds = metrics.pose_metrics(model, pa, pb)
    assert float(add[0]) == pytest.approx(0.01, rel=1e-5)
    assert float(adds[0]) <= float(add[0]) + 1e-9


def test_no_cpu_fallback():
    with pytest.raises(_cabi.Plr2Error):
        metrics.auc(torch.zeros(4))
    with pytest.raises(_cabi.Plr2Error):
        metrics.pose_metrics(torch.zeros(4, 3), torch.zeros(1, 12), torch.zeros(1, 12))


def test_rotation_translation_errors_match_reference_golden():
    g = load_golden("metrics_add_adds")
    rot, tr = metrics.pose_errors(_cuda(g["err_pose1"]), _cuda(g["err_pose2"]))
    # float32 rotation matrices: entries carry ~6e-8 error, so angles are good to ~3e-4 rad only where the two
    # rotations coincide (sqrt of the rounding) and to ~1e-6 elsewhere; rtol 1e-3 as everywhere
    np.testing.assert_allclose(rot.cpu().numpy(), g["err_rot"], rtol=RTOL, atol=5e-4)
    np.testing.assert_allclose(tr.cpu().numpy(), g["err_trans"], rtol=RTOL, atol=1e-6)
    assert abs(float(rot[2]) - np.pi) < 1e-3                    # the 180 degree case
