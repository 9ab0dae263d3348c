"""Generates tests/golden/kalman.npz from the REFERENCE's own pose Kalman filter (src/estimation/filter.py:107-202,
state.py, motion.py), imported from /root/reference/src.  Run in the build container only:
    python tests/golden/make_kalman_golden.py

Stored (all float64):
  comp_*   apply_transform / relative_transform of State_R3xQuat pairs (state.py:137-147), incl. rotations whose
           from_matrix branch is each of the four cases (quaternion sign convention)
  jac_*    the reference's apply_transform differenced with h = 1e-6 (its own code, state.py:150-161, with only the
           step changed from machine epsilon): pins the analytic Jacobian
  upd_*    Kalman_Filter.update outputs (state and covariance) for diagonal and full prior covariances
  prd_*    Kalman_Filter.predict: the predicted STATE (deterministic; its covariance uses the eps-step Jacobian and is
           not stored)
  seq_*    whole filtered trajectories of the loop of src/helper/postprocess.py:35-61, produced by the reference
           Kalman_Filter with State_R3xQuat.Jacobian_apply_transform replaced by the SAME function with h = 1e-6
"""
import os
import sys
import types

import numpy as np

try:
    import cv2  # noqa: F401
except Exception:                                    # pragma: no cover
    sys.modules["cv2"] = types.ModuleType("cv2")
sys.path.insert(0, "/root/reference/src")
from estimation.filter import Kalman_Filter       # noqa: E402
from estimation.state import State_R3xQuat        # noqa: E402
from scipy.spatial.transform import Rotation       # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
H = 1e-6


def jac_h(self, pose2, h=H):
    """state.py:150-161 verbatim except for the step size."""
    J = np.zeros((7, 7), dtype=np.float64)
    for i in range(0, 7):
        pos = State_R3xQuat(self.state.copy())
        pos.state[i, 0] += h
        neg = State_R3xQuat(self.state.copy())
        neg.state[i, 0] -= h
        col = np.true_divide((pos.apply_transform(pose2).state - neg.apply_transform(pose2).state), 2 * h)
        J[:, i] = np.squeeze(col)
    return J


def rand_pose(rng, ang=1.0, ts=0.5, unit=True):
    q = Rotation.from_rotvec(rng.normal(size=3) * ang).as_quat()
    if rng.random() < 0.5:
        q = -q
    if not unit:
        q = q * rng.uniform(0.7, 1.4)
    return np.concatenate([rng.normal(size=3) * ts, q])


def main():
    rng = np.random.default_rng(4711)
    out = {}
    # ---- composition
    n = 40
    p1 = np.stack([rand_pose(rng, ang=2.5) for _ in range(n)])
    p2 = np.stack([rand_pose(rng, ang=2.5) for _ in range(n)])
    # force all four from_matrix branches: rotations by ~pi about x, y, z and small rotations
    for i, ax in enumerate(([1, 0, 0], [0, 1, 0], [0, 0, 1])):
        p1[i, 3:] = Rotation.from_rotvec(np.array(ax) * 3.0).as_quat()
        p2[i, 3:] = Rotation.from_rotvec(rng.normal(size=3) * 0.05).as_quat()
    out["comp_p1"], out["comp_p2"] = p1, p2
    out["comp_apply"] = np.stack([State_R3xQuat(a).apply_transform(State_R3xQuat(b)).state[:, 0] for a, b in zip(p1, p2)])
    out["comp_rel"] = np.stack([State_R3xQuat(a).relative_transform(State_R3xQuat(b)).state[:, 0] for a, b in zip(p1, p2)])
    # ---- Jacobian (h = 1e-6), unit and non-unit quaternions in p1
    m = 16
    j1 = np.stack([rand_pose(rng, ang=1.5, unit=(i % 2 == 0)) for i in range(m)])
    j2 = np.stack([rand_pose(rng, ang=0.3, ts=0.05) for _ in range(m)])
    out["jac_p1"], out["jac_p2"] = j1, j2
    out["jac_J"] = np.stack([jac_h(State_R3xQuat(a), State_R3xQuat(b)) for a, b in zip(j1, j2)])
    # ---- update
    v7 = (0.1, 0.1, 0.1, 0.05, 0.05, 0.05, 0.05)
    k = 12
    us, uP, uo, ur, os_, oP = [], [], [], [], [], []
    for i in range(k):
        kf = Kalman_Filter(state_type="State_R3xQuat", motion_model="Linear_Motion", variance_motion=v7, variance_sensor=v7)
        prior = rand_pose(rng)
        if i % 2 == 0:
            P = np.diag(rng.uniform(0.01, 10, 7))
        else:
            A = rng.normal(size=(7, 7))
            P = A @ A.T * 0.1 + np.diag(rng.uniform(0.01, 1, 7))
        kf.set_prior(prior, np.diag(P).copy())
        kf.covariances = [P.copy()]                  # set_prior turns a 7x7 argument into its diagonal VECTOR (len() == 7, filter.py:154-155)
        obs = rand_pose(rng)
        r = rng.uniform(0.01, 2.0, 7)
        s, Pn = kf.update(State_R3xQuat(obs), r)
        us.append(prior); uP.append(P); uo.append(obs); ur.append(r); os_.append(s.state[:, 0].copy()); oP.append(np.array(Pn))
    out.update(upd_state=np.stack(us), upd_P=np.stack(uP), upd_obs=np.stack(uo), upd_r=np.stack(ur),
               upd_out_state=np.stack(os_), upd_out_P=np.stack(oP))
    # ---- predict (state only), a_t / a_r variants
    pa, pb, pp, par = [], [], [], []
    for i in range(12):
        a_t, a_r = [(1.0, 1.0), (0.5, 1.0), (1.0, 0.5), (0.3, 0.7)][i % 4]
        kf = Kalman_Filter(state_type="State_R3xQuat", motion_model="Linear_Motion", variance_motion=v7, variance_sensor=v7,
                           params={"a_t": a_t, "a_r": a_r})
        a = rand_pose(rng, ang=0.8)
        b = State_R3xQuat(a).apply_transform(State_R3xQuat(rand_pose(rng, ang=0.2, ts=0.05))).state[:, 0].copy()
        kf.set_prior(a, np.array(v7))
        kf.predictions.append(State_R3xQuat(b))
        kf.covariances.append(np.diag(v7))
        pred, _ = kf.predict()
        pa.append(a); pb.append(b); pp.append(pred.state[:, 0].copy()); par.append([a_t, a_r])
    out.update(prd_a=np.stack(pa), prd_b=np.stack(pb), prd_out=np.stack(pp), prd_params=np.array(par))
    # ---- whole sequences, reference filter with the h = 1e-6 Jacobian
    State_R3xQuat.Jacobian_apply_transform = jac_h
    S, T = 6, 20
    priors, obs_all, var_all, filt_all, pd_all, prd_all, params = [], [], [], [], [], [], []
    qd = np.array([0.02, 0.02, 0.02, 0.01, 0.01, 0.01, 0.01])
    sv = np.array([0.05, 0.05, 0.05, 0.02, 0.02, 0.02, 0.02])
    for s_i in range(S):
        a_t, a_r = [(1.0, 1.0), (1.0, 1.0), (0.5, 1.0), (1.0, 0.6), (0.8, 0.8), (1.0, 1.0)][s_i]
        kf = Kalman_Filter(state_type="State_R3xQuat", motion_model="Linear_Motion", variance_motion=tuple(qd),
                           variance_sensor=tuple(sv), params={"a_t": a_t, "a_r": a_r})
        prior = rand_pose(rng, ang=0.5)
        kf.set_prior(prior, np.array([10, 10, 10, 10, 10, 10, 20 * np.pi]))
        # a smooth trajectory + observation noise; confidences in (0.2, 1)
        base = rand_pose(rng, ang=0.8)
        step = rand_pose(rng, ang=0.06, ts=0.01)
        cur = State_R3xQuat(base)
        obs, var, filt, pd, prd = [], [], [], [], []
        for t in range(T):
            cur = cur.apply_transform(State_R3xQuat(step))
            o = cur.state[:, 0].copy()
            o[:3] += rng.normal(size=3) * 0.004
            o[3:] += rng.normal(size=4) * 0.004
            o[3:] /= np.linalg.norm(o[3:])
            if s_i == 1 and t % 5 == 0:
                o[3:] = -o[3:]                       # the detector may return either sign: linear arithmetic on q sees it
            c = rng.uniform(0.2, 1.0)
            v = np.true_divide(sv, c)
            f, Pn = kf.update(State_R3xQuat(o), v)
            filt.append(f.state[:, 0].copy()); pd.append(np.diag(Pn).copy())
            kf.predict()
            prd.append(kf.predictions[-1].state[:, 0].copy())
            obs.append(o); var.append(v)
        priors.append(prior); obs_all.append(obs); var_all.append(var); filt_all.append(filt); pd_all.append(pd); prd_all.append(prd)
        params.append([a_t, a_r])
    out.update(seq_prior=np.stack(priors), seq_obs=np.array(obs_all), seq_var=np.array(var_all), seq_q=qd,
               seq_prior_var=np.array([10, 10, 10, 10, 10, 10, 20 * np.pi]), seq_params=np.array(params),
               seq_filtered=np.array(filt_all), seq_pdiag=np.array(pd_all), seq_predicted=np.array(prd_all))
    path = os.path.join(HERE, "kalman.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
