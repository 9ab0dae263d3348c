"""GPU: the stand-alone tcgen05 split-bf16 GEMM entry point (plr2_gemm_bf16x2) against torch float64 over shapes
that exercise ragged M, partial and half-width last N tiles (N % 256 in (0, 128] runs the tail tile at half MMA
width), few-tile problems (narrow tiles) and K from one to sixteen 64-wide blocks."""
import pytest

from tools.gemm_check import run, run_tn

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(1, 1920, 1024), (512, 1920, 1024), (1000, 640, 384), (3000, 384, 128),
                                   (777, 130, 64), (2048, 1152, 256), (5000, 320, 64), (300, 1920, 384),
                                   (255, 256, 64), (257, 512, 192), (40000, 256, 640)])
def test_gemm_bf16x2_matches_float64(shape):
    assert run(*shape) == 0          # every element within 1e-4 * (1 + |ref|): split-bf16 operands, fp32 accumulate


def test_gemm_bf16x2_relu_epilogue():
    assert run(1000, 640, 384, relu=1) == 0


@pytest.mark.parametrize("shape", [(128, 64, 64), (256, 256, 128), (64, 3, 1000), (1024, 512, 8000), (130, 200, 777),
                                   (1920, 384, 8000), (640, 256, 4100), (8, 128, 1000), (1920, 1024, 8)])
def test_gemm_bf16x2_tn_matches_float64(shape):
    """Weight-gradient form on MN-major planes (no transposed copy): ragged M / N / K, NaN-filled padding."""
    assert run_tn(*shape) == 0
