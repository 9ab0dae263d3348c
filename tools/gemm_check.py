#!/usr/bin/env python
"""Stand-alone check of plr2_gemm_bf16x2 (tcgen05 split-bf16 GEMM, FP32 output) against torch fp64 for a list of
shapes given as M,N,K triples.  Usage: python tools/gemm_check.py 1,1920,1024 512,1920,1024 ..."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from plr2_b200 import _cabi


def run(M, N, K, relu=0):
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N * 3 + K)
    A = torch.randn(M, K, device="cuda", generator=g)
    W = torch.randn(N, K, device="cuda", generator=g) / K ** 0.5
    b = torch.randn(N, device="cuda", generator=g)
    L = _cabi.lib()
    Ap = torch.empty(2, M, K, device="cuda", dtype=torch.bfloat16)
    Wp = torch.empty(2, N, K, device="cuda", dtype=torch.bfloat16)
    st = _cabi.stream_ptr(A.device)
    _cabi.check(L.plr2_split_bf16x2(_cabi.ptr(A), K, M, K, _cabi.ptr(Ap), K, M * K, st), "split A")
    _cabi.check(L.plr2_split_bf16x2(_cabi.ptr(W), K, N, K, _cabi.ptr(Wp), K, N * K, st), "split W")
    C = torch.full((M, N), float("nan"), device="cuda")
    _cabi.check(L.plr2_gemm_bf16x2(_cabi.ptr(Ap), K, M * K, _cabi.ptr(Wp), K, N * K, M, N, K, _cabi.ptr(b), relu,
                                   _cabi.ptr(C), N, st), "gemm")
    torch.cuda.synchronize()
    ref = A.double() @ W.double().T + b.double()
    if relu:
        ref = ref.clamp_min(0)
    err = (C.double() - ref).abs()
    bad = ~(err <= 1e-4 * (1 + ref.abs()))
    print(f"M={M} N={N} K={K}: max err {err.nan_to_num(9e9).max().item():.3e}, bad {int(bad.sum())}/{M * N}, nan {int(C.isnan().sum())}")
    if bad.any():
        idx = bad.nonzero()
        print("  first bad (row, col):", idx[:5].tolist(), " bad rows:", sorted(set(idx[:, 0].tolist()))[:10],
              " bad col range:", int(idx[:, 1].min()), int(idx[:, 1].max()))
    return int(bad.sum())


if __name__ == "__main__":
    shapes = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]] or [(1, 1920, 1024), (512, 1920, 1024)]
    sys.exit(1 if sum(run(*s) for s in shapes) else 0)
