"""ORACLE - TEST INFRASTRUCTURE ONLY (never imported by plr2_b200/).

numpy float64 restatement of the reference's pose Kalman filter:

    quat_to_matrix / matrix_to_quat  <- scipy Rotation.from_quat().as_matrix() / from_matrix().as_quat() as used by
                                        src/estimation/state.py:58,169 (xyzw order, NO sign canonicalisation:
                                        the branch on the largest of m00, m11, m22, trace decides the sign)
    apply_transform                  <- State_R3xQuat.apply_transform   src/estimation/state.py:137-139
    relative_transform               <- State_R3xQuat.relative_transform :145-147
    interpolate_motion               <- Linear_Motion.interpolate_motion src/estimation/motion.py:27-47
    kalman_update                    <- Kalman_Filter.update            src/estimation/filter.py:159-189
                                        (incl. its ELEMENT-WISE covariance update np.multiply((I - K o H), P), :186,
                                        which leaves P diagonal after every update)
    kalman_predict                   <- Kalman_Filter.predict           src/estimation/filter.py:132-146
    run_sequence                     <- the loop of kf_sequence         src/helper/postprocess.py:35-61

ONE stated deviation: the reference propagates the covariance with a central-difference Jacobian whose step is
h = np.finfo(float64).eps (state.py:150-161), i.e. rounding noise divided by 4.4e-16 (profiles/r2_kalman_reference_probe.txt).
`jacobian_apply_transform` here is the ANALYTIC derivative of the same composition.  Pinned by
tests/golden/make_kalman_golden.py on the reference's own classes: update(), the predicted states, the composition
functions, the analytic Jacobian against the reference's apply_transform differenced with h = 1e-6, and whole filtered
trajectories against the reference Kalman_Filter run with ONLY that step size changed.
"""
import numpy as np


def quat_to_matrix(q):
    """q (4,) xyzw, any norm (scipy normalises in from_quat) -> (3,3)."""
    q = np.asarray(q, dtype=np.float64)
    x, y, z, w = q / np.linalg.norm(q)
    x2, y2, z2, w2 = x * x, y * y, z * z, w * w
    xy, zw, xz, yw, yz, xw = x * y, z * w, x * z, y * w, y * z, x * w
    return np.array([[x2 - y2 - z2 + w2, 2 * (xy - zw), 2 * (xz + yw)],
                     [2 * (xy + zw), -x2 + y2 - z2 + w2, 2 * (yz - xw)],
                     [2 * (xz - yw), 2 * (yz + xw), -x2 - y2 + z2 + w2]])


def matrix_to_quat(m):
    """(3,3) proper orthogonal -> xyzw unit quaternion with scipy's sign convention."""
    tr = m[0, 0] + m[1, 1] + m[2, 2]
    dec = [m[0, 0], m[1, 1], m[2, 2], tr]
    c = int(np.argmax(dec))
    q = np.empty(4)
    if c != 3:
        i, j, k = c, (c + 1) % 3, (c + 2) % 3
        q[i] = 1 - tr + 2 * m[i, i]
        q[j] = m[j, i] + m[i, j]
        q[k] = m[k, i] + m[i, k]
        q[3] = m[k, j] - m[j, k]
    else:
        q[0] = m[2, 1] - m[1, 2]
        q[1] = m[0, 2] - m[2, 0]
        q[2] = m[1, 0] - m[0, 1]
        q[3] = 1 + tr
    return q / np.linalg.norm(q)


def apply_transform(p1, p2):
    """State (7,) = t xyz | q xyzw.  SE3(p1) @ SE3(p2) converted back (state.py:137-139)."""
    r1, r2 = quat_to_matrix(p1[3:]), quat_to_matrix(p2[3:])
    out = np.empty(7)
    out[:3] = r1 @ p2[:3] + p1[:3]
    out[3:] = matrix_to_quat(r1 @ r2)
    return out


def inv_pose(p):
    r = quat_to_matrix(p[3:])
    out = np.empty(7)
    out[:3] = -r.T @ p[:3]
    out[3:] = matrix_to_quat(r.T)
    return out


def relative_transform(p1, p2):
    return apply_transform(inv_pose(p1), p2)


IDENTITY = np.array([0, 0, 0, 0, 0, 0, 1.0])


def interpolate_motion(traj, a_t=1.0, a_r=1.0):
    """Motion between the last two states of `traj`, scaled by a_t / slerped by a_r (motion.py:27-47; t = -1 there).
    Fewer than two states: the identity motion.  Identical consecutive states give theta = 0 and 0/0 = NaN in the
    reference's slerp; that is reproduced."""
    if len(traj) < 2:
        return IDENTITY.copy()
    m = relative_transform(traj[-2], traj[-1])
    m[:3] = (1 - a_t) * IDENTITY[:3] + a_t * m[:3]
    with np.errstate(invalid="ignore", divide="ignore"):
        theta = np.arccos(np.dot(IDENTITY[3:], m[3:]))
        m[3:] = np.true_divide(np.sin((1 - a_r) * theta) * IDENTITY[3:] + np.sin(a_r * theta) * m[3:], np.sin(theta))
    return m


def jacobian_apply_transform(p1, p2):
    """d apply_transform(p1, p2) / d p1 (7,7), analytic - the derivative state.py:150-161 differences with h = eps.
    t = t1 + R(q1/|q1|) t2;  q = s * (q1/|q1|) (x) (q2/|q2|), s = +-1 the sign matrix_to_quat picks."""
    q1 = np.asarray(p1[3:], dtype=np.float64)
    n1 = np.linalg.norm(q1)
    u = q1 / n1
    proj = (np.eye(4) - np.outer(u, u)) / n1                    # d u / d q1
    x, y, z, w = u
    t2 = np.asarray(p2[:3], dtype=np.float64)
    # d (R(u) v) / d u for a UNIT-parametrised homogeneous form R = as_matrix above (valid on |u| = 1 after projection)
    def dR(v):
        a, b, c = v
        return 2 * np.array([[x * a + y * b + z * c, -y * a + x * b + w * c, -z * a - w * b + x * c, w * a - z * b + y * c],
                             [y * a - x * b - w * c, x * a + y * b + z * c, w * a - z * b + y * c, z * a + w * b - x * c],
                             [z * a + w * b - x * c, -w * a + z * b - y * c, x * a + y * b + z * c, -y * a + x * b + w * c]])
    J = np.zeros((7, 7))
    J[:3, :3] = np.eye(3)
    J[:3, 3:] = dR(t2) @ proj
    v = np.asarray(p2[3:], dtype=np.float64)
    v = v / np.linalg.norm(v)
    bx, by, bz, bw = v
    # Hamilton product u (x) v (xyzw) is linear in u: prod = Rmat(v) u
    right = np.array([[bw, bz, -by, bx],
                      [-bz, bw, bx, by],
                      [by, -bx, bw, bz],
                      [-bx, -by, -bz, bw]])
    prod = right @ u
    q_out = apply_transform(p1, p2)[3:]
    s = 1.0 if np.dot(q_out, prod) >= 0 else -1.0
    J[3:, 3:] = s * right @ proj
    return J


def kalman_update(state, P, obs, r_diag):
    """filter.py:159-189 with H = I.  Returns (state', P')."""
    R = np.diag(np.asarray(r_diag, dtype=np.float64))
    K = P @ np.linalg.inv(P + R)
    new = state + K @ (obs - state)
    new[3:] = new[3:] / np.linalg.norm(new[3:])
    Pn = np.multiply(np.identity(7) - np.multiply(K, np.identity(7)), P)     # element-wise, as the reference does
    return new, Pn


def kalman_predict(traj, P, q_diag, a_t=1.0, a_r=1.0, jacobian=jacobian_apply_transform):
    """filter.py:132-146 (t = -1): returns (predicted state, predicted covariance)."""
    m = interpolate_motion(traj, a_t, a_r)
    pred = apply_transform(traj[-1], m)
    pred[3:] = pred[3:] / np.linalg.norm(pred[3:])
    A = jacobian(traj[-1], m)
    return pred, A @ P @ A.T + np.diag(np.asarray(q_diag, dtype=np.float64))


def run_sequence(prior, prior_var, obs, sensor_var, q_diag, a_t=1.0, a_r=1.0, jacobian=jacobian_apply_transform):
    """postprocess.py:35-61 for one sequence: set_prior, then per observation update -> predict.
    obs (T,7), sensor_var (T,7) (= var_sensor / obs_c).  Returns filtered (T,7), p_diag (T,7) = diag of P after each
    update, predicted (T,7) = the state predict() appended after each step."""
    traj = [np.asarray(prior, dtype=np.float64).copy()]
    P = np.diag(np.asarray(prior_var, dtype=np.float64))
    T = len(obs)
    filt, pd, prd = np.empty((T, 7)), np.empty((T, 7)), np.empty((T, 7))
    for t in range(T):
        traj[-1], P = kalman_update(traj[-1], P, np.asarray(obs[t], dtype=np.float64), sensor_var[t])
        filt[t], pd[t] = traj[-1], np.diag(P)
        nxt, P = kalman_predict(traj, P, q_diag, a_t, a_r, jacobian)
        traj.append(nxt)
        prd[t] = nxt
    return filt, pd, prd
