"""CPU: the error model behind the split-bf16 contractions and the sign-safe ReLU pass (csrc/gemm_tc_f32.cu).

x = hi + lo + e with hi = bf16(x), lo = bf16(x - hi): |e| <= 2^-16 |x|.  The tensor-core product uses
hi*hi + hi*lo + lo*hi (the lo*lo term is dropped), so per term the relative error is at most 3 * 2^-16 and over a
dot product |error| <= 3 * 2^-16 * sum |x_k w_k| <= 3 * 2^-16 * ||x|| ||w||  <  SIGN_TAU * ||x|| ||w||, SIGN_TAU = 2^-14."""
import numpy as np
import torch

SIGN_TAU = 2.0 ** -14


def split(v):
    hi = v.to(torch.bfloat16).to(torch.float32)
    lo = (v - hi).to(torch.bfloat16).to(torch.float32)
    return hi, lo


def test_split_residual_is_below_2_pow_minus_16():
    g = torch.Generator().manual_seed(0)
    v = torch.randn(1 << 16, generator=g) * torch.exp(4 * torch.randn(1 << 16, generator=g))
    hi, lo = split(v)
    rel = ((v.double() - hi.double() - lo.double()).abs() / v.double().abs()).max().item()
    assert rel <= 2.0 ** -16


def test_three_product_error_is_inside_the_sign_safe_bound():
    g = torch.Generator().manual_seed(1)
    worst = 0.0
    for k in (64, 384, 512, 640, 1024):
        x = torch.relu(torch.randn(256, k, generator=g))            # post-ReLU activations
        w = torch.randn(128, k, generator=g) / np.sqrt(k)
        xh, xl = split(x)
        wh, wl = split(w)
        tc = (xh.double() @ wh.double().T) + (xh.double() @ wl.double().T) + (xl.double() @ wh.double().T)
        exact = x.double() @ w.double().T
        bound = SIGN_TAU * x.double().norm(dim=1)[:, None] * w.double().norm(dim=1)[None, :]
        ratio = ((tc - exact).abs() / bound).max().item()
        worst = max(worst, ratio)
        # every element whose sign could differ lies inside the bound, and few elements do
        risky = (tc.sign() != exact.sign())
        assert not (risky & ((tc.abs() >= bound))).any()
        assert (tc.abs() < bound).double().mean().item() < 0.01
    assert worst < 0.25        # typical errors sit far inside the rigorous bound
