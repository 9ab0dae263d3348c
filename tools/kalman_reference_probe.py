#!/usr/bin/env python
"""Evidence for SURVEY.md 8(f) row 4 (batched pose Kalman filter, src/estimation/filter.py:107-202): why the device filter
(csrc/kalman.cu) propagates its covariance with the ANALYTIC Jacobian instead of reproducing the reference's.  Run in
the build container (imports the reference from /root/reference/src):

    python tools/kalman_reference_probe.py > profiles/r2_kalman_reference_probe.txt

The reference propagates its covariance with `State_R3xQuat.Jacobian_apply_transform` (src/estimation/state.py:150-161):
a central difference with step h = np.finfo(np.float64).eps.  At state values of order 1, x +- h moves x by one or two
units in the last place (or not at all), so (f(x+h) - f(x-h)) / 2h is the ROUNDING NOISE of scipy's Rotation
(from_quat -> as_matrix -> matmul -> from_matrix -> as_quat) and of numpy's 4x4 matmul, divided by 4.4e-16: entries
are multiples of 1/4, 1/2 ... that have nothing to do with the derivative and change with the BLAS / scipy build.  The
Kalman gain K = P (P + R)^-1 and hence every filtered pose inherit that noise, so bit-level agreement with it has no
build-independent meaning.  Everything else of the filter is deterministic and is pinned on the reference's own
outputs (tests/golden/make_kalman_golden.py, tests/test_kalman.py), whole trajectories included - against the
reference class run with only that step size changed to 1e-6.
"""
import sys
import types

import numpy as np

try:
    import cv2  # noqa: F401
except Exception:                                    # pragma: no cover
    sys.modules["cv2"] = types.ModuleType("cv2")
sys.path.insert(0, "/root/reference/src")
from estimation.filter import Kalman_Filter      # noqa: E402
from estimation.state import State_R3xQuat       # noqa: E402
from scipy.spatial.transform import Rotation      # noqa: E402


def analytic_jacobian(p1, p2, h=1e-6):
    """The derivative the reference MEANT: same composition, step 1e-6."""
    J = np.zeros((7, 7))
    for i in range(7):
        a, b = State_R3xQuat(p1.state.copy()), State_R3xQuat(p1.state.copy())
        a.state[i, 0] += h
        b.state[i, 0] -= h
        J[:, i] = np.squeeze((a.apply_transform(p2).state - b.apply_transform(p2).state) / (2 * h))
    return J


def main():
    np.set_printoptions(precision=3, suppress=True, linewidth=160)
    rng = np.random.default_rng(0)
    q1 = Rotation.from_rotvec(rng.normal(size=3) * 0.4).as_quat()
    q2 = Rotation.from_rotvec(rng.normal(size=3) * 0.05).as_quat()
    p1 = State_R3xQuat(np.concatenate([[0.1, -0.2, 0.9], q1]))
    p2 = State_R3xQuat(np.concatenate([[0.01, 0.0, -0.02], q2]))
    Jr = p1.Jacobian_apply_transform(p2)
    Ja = analytic_jacobian(p1, p2)
    print("reference Jacobian_apply_transform (h = eps = %.3g):" % np.finfo(np.float64).eps)
    print(Jr)
    print("the same derivative with h = 1e-6:")
    print(Ja)
    print("max |difference| = %.3f   (entries of the reference matrix that are integer multiples of 1/8: %d of 49)"
          % (np.abs(Jr - Ja).max(), int(np.sum(np.abs(Jr * 8 - np.round(Jr * 8)) < 1e-9))))
    # sensitivity of the filter output to a perturbation far below any tolerance
    def run(eps_shift):
        traj = []
        x = np.array([0.0, 0.0, 1.0])
        for t in range(12):
            q = Rotation.from_rotvec([0.02 * t, 0.01 * t, 0.03 * t]).as_quat()
            traj.append(State_R3xQuat(np.concatenate([x + [0.01 * t, 0.0, 0.005 * t], q])))
        traj[3].state[0, 0] += eps_shift
        v7 = (0.1, 0.1, 0.1, 0.05, 0.05, 0.05, 0.05)    # the class defaults have 8 entries and fail to broadcast (filter.py:113-114)
        kf = Kalman_Filter(state_type="State_R3xQuat", motion_model="Linear_Motion", variance_motion=v7, variance_sensor=v7)
        kf.set_prior(traj[0], np.array([10, 10, 10, 10, 10, 10, 20 * np.pi]))
        out = []
        for o in traj[1:]:
            f, _ = kf.update(o, np.array([0.1, 0.1, 0.1, 0.05, 0.05, 0.05, 0.05]))
            kf.predict()
            out.append(f.state[:, 0].copy())
        return np.array(out)
    a, b = run(0.0), run(1e-13)
    print("filtered trajectory, 11 steps; one observation shifted by 1e-13 m:")
    print("max |change of a filtered state component| = %.3e  (input change 1e-13)" % np.abs(a - b).max())


if __name__ == "__main__":
    main()
