"""CPU: the operand-format sweep behind the precision scheme of the tcgen05 GEMM chain (tools/precision_sweep.py,
profiles/r2_precision_sweep_cpu.txt), asserted on a reduced sample: which formats hold the north-star criterion
(>= 95 % of the per-point outputs within rtol 1e-3 of a float64 evaluation) and which do not.  The products of two
<= 11-bit operands are exact in FP32, so rounding the operands and multiplying in FP32 reproduces the tensor-core
arithmetic up to the summation order."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import precision_sweep as P  # noqa: E402
from plr2_b200 import synth  # noqa: E402


@pytest.fixture(scope="module")
def setup():
    torch.set_num_threads(min(8, os.cpu_count() or 1))
    shape = synth.YCB
    params = {k: torch.from_numpy(v) for k, v in synth.make_params("posenet", shape.num_obj, 1).items()}
    frames = [synth.make_frame(shape, 3000 + 17 * i, obj=o) for i, o in enumerate([12, 3])]
    return shape, params, frames


def _min_frac(setup, fmt):
    shape, params, frames = setup
    e = P.evaluate(fmt, frames, params, shape, "cpu")
    return min(e["r"][0], e["t"][0], e["c"][0])


def test_shipped_scheme_holds_with_margin(setup):
    shipped = dict({g: "fp16x3" for g in P.GROUPS}, c5="fp16x2", c6="fp16x2")      # csrc/posenet.cu, PLR2_INFER_FMT default
    round1 = P.ROUND1                                                                # bf16 planes (PLR2_INFER_FMT=bf16)
    a, b = _min_frac(setup, shipped), _min_frac(setup, round1)
    assert a >= 0.99 and b >= 0.97 and a >= b          # fp16 hi/lo planes: same MMA count, more mantissa


@pytest.mark.parametrize("fmt", ["bf16x1", "fp16x1", "tf32x1", "bf16x2", "fp16x2"])
def test_single_rounded_operands_fail_on_the_heads(setup, fmt):
    """One rounding of either operand of a per-point head layer (bf16 8 bits, fp16 / TF32 11 bits) breaks the criterion:
    the 3-MMA split is not optional there.  kind::tf32 would cost 2 bf16-rate MMAs per MAC and fails like fp16 x 1."""
    base = {g: "fp16x3" for g in P.GROUPS}
    for layer in ("h1", "h2", "h3"):
        assert _min_frac(setup, dict(base, **{layer: fmt})) < 0.95, (layer, fmt)


def test_everything_single_product_fails(setup):
    for fmt in ("bf16x1", "fp16x1", "tf32x1"):
        assert _min_frac(setup, {g: fmt for g in P.GROUPS}) < 0.90


def test_refiner_single_product_conv6_costs_20x_accuracy():
    """PoseRefineNet sees conv6 only through the 1000-point mean: activation rounding averages out, weight rounding does
    not.  X_hi (W_hi + W_lo) in fp16 is ~10x closer to float64 than round 1's bf16 form; X_hi W_hi is ~20x further."""
    lines = []
    P.sweep_refiner(4, param_seeds=(2,), log=lines.append)
    err = {}
    for l in lines:
        name = l[len("refiner params 2  "):].split(" max |err|")[0].strip()
        err[name] = max(float(l.split("pred_r")[1].split()[0]), float(l.split("pred_t")[1].split()[0]))
    assert err["fp16; c2,c5,c6 = x2"] < 2e-5 < 5e-5 < err["fp16; c2,c5 = x2, c6 = x1"]
    assert err["fp16; c2,c5,c6 = x2"] < err["round-1 scheme (bf16; c2,c5,c6 = x2)"]
