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


def run_tn(M, N, K):
    """plr2_gemm_bf16x2_tn: C[M][N] = A^T B, A [K][M], B [K][N] as split planes with padded pitches; K arbitrary."""
    g = torch.Generator(device="cuda").manual_seed(M * 5 + N * 11 + K)
    A = torch.randn(K, M, device="cuda", generator=g)
    B = torch.randn(K, N, device="cuda", generator=g) / K ** 0.5
    L = _cabi.lib()
    lda, ldb = (M + 63) // 64 * 64, (N + 7) // 8 * 8
    # NaN in the padding columns and in rows past K: neither may reach the result
    Ap = torch.full((2, K + 3, lda), float("nan"), device="cuda", dtype=torch.bfloat16)
    Bp = torch.full((2, K + 3, ldb), float("nan"), device="cuda", dtype=torch.bfloat16)
    st = _cabi.stream_ptr(A.device)
    _cabi.check(L.plr2_split_bf16x2(_cabi.ptr(A), M, K, M, _cabi.ptr(Ap), lda, (K + 3) * lda, st), "split A")
    _cabi.check(L.plr2_split_bf16x2(_cabi.ptr(B), N, K, N, _cabi.ptr(Bp), ldb, (K + 3) * ldb, st), "split B")
    C = torch.full((M, N), float("nan"), device="cuda")
    _cabi.check(L.plr2_gemm_bf16x2_tn(_cabi.ptr(Ap), lda, (K + 3) * lda, _cabi.ptr(Bp), ldb, (K + 3) * ldb, M, N, K,
                                      _cabi.ptr(C), N, st), "gemm_tn")
    torch.cuda.synchronize()
    ref = A.double().T @ B.double()
    err = (C.double() - ref).abs()
    bad = ~(err <= 1e-4 * (1 + ref.abs()))
    print(f"TN M={M} N={N} K={K}: max err {err.nan_to_num(9e9).max().item():.3e}, bad {int(bad.sum())}/{M * N}, nan {int(C.isnan().sum())}")
    if bad.any():
        idx = bad.nonzero()
        print("  first bad (row, col):", idx[:5].tolist(), " bad rows:", sorted(set(idx[:, 0].tolist()))[:10],
              " bad col range:", int(idx[:, 1].min()), int(idx[:, 1].max()))
    return int(bad.sum())


if __name__ == "__main__":
    if sys.argv[1:2] == ["tn"]:
        shapes = [tuple(int(v) for v in a.split(",")) for a in sys.argv[2:]] or [(128, 64, 64), (256, 256, 128), (1920, 384, 8000)]
        sys.exit(1 if sum(run_tn(*s) for s in shapes) else 0)
    shapes = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]] or [(1, 1920, 1024), (512, 1920, 1024)]
    sys.exit(1 if sum(run(*s) for s in shapes) else 0)
