"""Shared helpers for the parity tests."""
import os

import numpy as np
import torch

from plr2_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

RTOL = 1e-3          # north_star: losses, poses and gradients within rtol 1e-3
FRAC = 0.95          # ... for >= 95 % of compared elements


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz")))


def shape_of(g):
    n, o, m, h, w = [int(v) for v in g["shape"]]
    return synth.Shape(n, o, m, h, w, tuple(int(v) for v in g["sym_list"]))


def frac_close(a, b, rtol=RTOL, atol=0.0):
    a = np.asarray(a, dtype=np.float64).reshape(-1)
    b = np.asarray(b, dtype=np.float64).reshape(-1)
    assert a.shape == b.shape, (a.shape, b.shape)
    return float(np.mean(np.abs(a - b) <= atol + rtol * np.abs(b)))


def assert_close(a, b, what, rtol=RTOL, atol=None, frac=FRAC):
    """>= frac of the elements within rtol (plus an absolute floor scaled to the tensor: gradients
    and coordinates have entries arbitrarily close to zero)."""
    b = np.asarray(b)
    if atol is None:
        atol = rtol * float(np.abs(b).max()) * 1e-1 if b.size else 0.0
    f = frac_close(a, b, rtol, atol)
    assert f >= frac, f"{what}: only {f:.4f} of elements within rtol={rtol} atol={atol:.3g}"
    return f


def tt(a, device="cpu", grad=False):
    x = torch.from_numpy(np.ascontiguousarray(a)).to(device)
    if grad:
        x.requires_grad_(True)
    return x


def knn_gap_ok(idx_test, idx_ref, best, second, gap=1e-6):
    """north_star rule: indices must agree wherever best and second-best differ by more than gap."""
    idx_test = np.asarray(idx_test).reshape(-1)
    idx_ref = np.asarray(idx_ref).reshape(-1)
    decided = (np.asarray(second).reshape(-1) - np.asarray(best).reshape(-1)) > gap
    bad = decided & (idx_test != idx_ref)
    return int(bad.sum()), int(decided.sum()), int((idx_test != idx_ref).sum())
