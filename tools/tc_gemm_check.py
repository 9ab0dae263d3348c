"""Stand-alone check of the tcgen05 split-bf16 GEMM against float64 numpy."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from plr2_b200 import _cabi  # noqa: E402


def split(x):
    rows, cols = x.shape
    planes = torch.empty((2, rows, cols), dtype=torch.bfloat16, device="cuda")
    rc = _cabi.lib().plr2_split_bf16x2(x.data_ptr(), cols, rows, cols, planes.data_ptr(), cols, rows * cols,
                                       torch.cuda.current_stream().cuda_stream)
    _cabi.check(rc, "split")
    return planes


def run(M, N, K, relu=False, iters=0):
    g = torch.Generator(device="cuda").manual_seed(M + 3 * N + 7 * K)
    X = torch.randn(M, K, device="cuda", generator=g)
    W = torch.randn(N, K, device="cuda", generator=g) / np.sqrt(K)
    b = torch.randn(N, device="cuda", generator=g)
    Xp, Wp = split(X), split(W)
    C = torch.full((M, N), float("nan"), device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    L = _cabi.lib()
    rc = L.plr2_gemm_bf16x2(Xp.data_ptr(), K, M * K, Wp.data_ptr(), K, N * K, M, N, K, b.data_ptr(), int(relu),
                            C.data_ptr(), N, st)
    _cabi.check(rc, "gemm_bf16x2")
    torch.cuda.synchronize()
    want = X.double() @ W.double().t() + b.double()
    if relu:
        want = want.clamp_min(0)
    err = (C.double() - want).abs().max().item()
    scale = want.abs().max().item()
    nan = int(torch.isnan(C).sum())
    msg = f"M={M} N={N} K={K} relu={relu}: max_abs_err={err:.3e} (scale {scale:.2f}) nan={nan}"
    if iters:
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(iters):
            L.plr2_gemm_bf16x2(Xp.data_ptr(), K, M * K, Wp.data_ptr(), K, N * K, M, N, K, b.data_ptr(), int(relu),
                               C.data_ptr(), N, st)
        ev1.record()
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1) / iters
        msg += f"  {ms*1e3:.1f} us  {2.0*M*N*K/ms/1e9:.1f} TFLOP/s (fp32-equivalent)"
    print(msg, flush=True)
    return err < 2e-4 * max(scale, 1.0) and nan == 0


if __name__ == "__main__":
    ok = True
    for shp in [(128, 128, 64), (128, 256, 64), (256, 128, 128), (200, 96, 64), (1000, 640, 384), (333, 1920, 384),
                (4096, 1024, 512), (1000, 256, 640), (130, 130, 192)]:
        ok &= run(*shp)
    ok &= run(1000, 512, 256, relu=True)
    for shp in [(512000, 1920, 384), (512000, 1024, 512), (512000, 256, 640), (64000, 1920, 384)]:
        ok &= run(*shp, relu=True, iters=5)
    print("ALL OK" if ok else "FAILED")
    sys.exit(0 if ok else 1)
