#!/usr/bin/env python
"""Timing of the batched pose Kalman filter (plr2_kalman_filter_f64) next to the numpy oracle on the host."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import kalman_oracle as KO  # noqa: E402
from plr2_b200 import estimation  # noqa: E402

rng = np.random.default_rng(0)
vm = np.array([0.02] * 3 + [0.01] * 4)
vs = np.array([0.05] * 3 + [0.02] * 4)
for S, T in ((64, 30), (4096, 30), (65536, 30)):
    q = rng.normal(size=(S, T, 4)) * 0.02 + np.array([0, 0, 0, 1.0])
    q /= np.linalg.norm(q, axis=-1, keepdims=True)
    obs = np.concatenate([np.cumsum(rng.normal(size=(S, T, 3)) * 0.01, 1) + [0, 0, 1], q], -1)
    conf = rng.uniform(0.2, 1, (S, T))
    o, c = torch.from_numpy(obs).cuda(), torch.from_numpy(conf).cuda()
    for _ in range(3):
        estimation.kalman_filter_sequences(o, c, vm, vs)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        f, _, _ = estimation.kalman_filter_sequences(o, c, vm, vs)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    n = min(S, 64)
    t0 = time.perf_counter()
    for s in range(n):
        KO.run_sequence([0, 0, 1, 0, 0, 0, 1.0], [10] * 6 + [20 * np.pi], obs[s], vs / conf[s][:, None], vm)
    cpu = (time.perf_counter() - t0) / n * S * 1e3
    print(f"S={S:6d} T={T}: device {ms:8.3f} ms = {S * T / ms / 1e3:9.1f} k filter steps/ms ... numpy oracle (1 core, extrapolated from {n} sequences) {cpu:10.1f} ms  ({cpu / ms:8.0f}x)",
          flush=True)
