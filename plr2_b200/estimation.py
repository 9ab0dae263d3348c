"""Batched pose Kalman filter on the device - the loop of src/helper/postprocess.py:35-61 over the
reference's Kalman_Filter (src/estimation/filter.py:107-202) for many sequences at once, through the C-ABI
(plr2_kalman_filter_f64).  float64 like the reference; CPU tensors raise Plr2Error (no host fallback).

    filtered, p_diag, predicted = kalman_filter_sequences(obs, obs_conf, var_motion, var_sensor,
                                                          prior=None, prior_var=None, params=None, lengths=None)

obs (S,T,7) = t xyz | q xyzw per observation (kf_sequence builds it from final_pred_obs['t'] and the wxyz quaternion
re-ordered by re_quat, postprocess.py:47-52), obs_conf (S,T) the network confidences: the sensor variance of an
observation is var_sensor / obs_c (postprocess.py:53).  prior / prior_var default to kf_sequence's
(0,0,1 | identity) and (10,10,10,10,10,10,20 pi) (postprocess.py:31-34; the reference adds random noise to its prior,
which the caller may do before passing it).  params: {'a_t': ..., 'a_r': ...} of Linear_Motion (motion.py:18-26).
"""
import math

import torch

from . import _cabi


def _cuda64(t, what, dev=None):
    if not isinstance(t, torch.Tensor):
        t = torch.as_tensor(t, dtype=torch.float64, device=dev)
    if not t.is_cuda:
        raise _cabi.Plr2Error("estimation.%s: CUDA tensors required (no CPU fallback in plr2_b200)" % what)
    return t.to(torch.float64).contiguous()


def kalman_filter_sequences(obs, obs_conf, var_motion, var_sensor, prior=None, prior_var=None, params=None, lengths=None):
    obs = _cuda64(obs, "kalman_filter_sequences")
    if obs.dim() != 3 or obs.shape[2] != 7:
        raise ValueError("estimation.kalman_filter_sequences: obs must be (S,T,7)")
    S, T, _ = obs.shape
    dev = obs.device
    vs = _cuda64(var_sensor, "kalman_filter_sequences", dev).reshape(7)
    if obs_conf is None:
        sensor_var = vs.expand(S, T, 7).contiguous()
    else:
        conf = _cuda64(obs_conf, "kalman_filter_sequences", dev).reshape(S, T, 1)
        sensor_var = torch.true_divide(vs, conf).contiguous()          # np.true_divide(var_sensor, obs_c), postprocess.py:53
    vm = _cuda64(var_motion, "kalman_filter_sequences", dev).reshape(7)
    if prior is None:
        prior = torch.tensor([0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0], dtype=torch.float64, device=dev).expand(S, 7)
    if prior_var is None:
        prior_var = torch.tensor([10.0] * 6 + [20 * math.pi], dtype=torch.float64, device=dev).expand(S, 7)
    prior = _cuda64(prior, "kalman_filter_sequences", dev).expand(S, 7).contiguous()
    prior_var = _cuda64(prior_var, "kalman_filter_sequences", dev).expand(S, 7).contiguous()
    a_t = float((params or {}).get("a_t", 1.0))
    a_r = float((params or {}).get("a_r", 1.0))
    ln = None
    if lengths is not None:
        ln = torch.as_tensor(lengths, device=dev).to(torch.int32).contiguous()
        if tuple(ln.shape) != (S,):
            raise ValueError("estimation.kalman_filter_sequences: lengths must be (S,)")
    filtered = torch.zeros(S, T, 7, dtype=torch.float64, device=dev)
    p_diag = torch.zeros(S, T, 7, dtype=torch.float64, device=dev)
    predicted = torch.zeros(S, T, 7, dtype=torch.float64, device=dev)
    if S == 0 or T == 0:
        return filtered, p_diag, predicted
    with torch.cuda.device(dev):
        rc = _cabi.lib().plr2_kalman_filter_f64(_cabi.ptr(prior), _cabi.ptr(prior_var), _cabi.ptr(obs), _cabi.ptr(sensor_var),
                                                _cabi.ptr(vm), _cabi.ptr(ln), a_t, a_r, S, T, _cabi.ptr(filtered),
                                                _cabi.ptr(p_diag), _cabi.ptr(predicted), _cabi.stream_ptr(dev))
    _cabi.check(rc, "plr2_kalman_filter_f64")
    return filtered, p_diag, predicted
