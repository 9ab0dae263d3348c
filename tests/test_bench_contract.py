"""CPU: the measurement contract of bench.py that can be checked without a GPU - the reference arm prints one
JSON line with the agreed keys, and our arm refuses to run (no CPU fallback) when there is no CUDA device."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(*args):
    env = dict(os.environ, OMP_NUM_THREADS="4")
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT, env=env,
                          capture_output=True, text=True, timeout=600)


def test_reference_arm_line():
    p = run("--impl", "reference", "--steps", "1", "--warmup", "0", "--ref-frames", "1")
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["unit"] == "poses/s" and j["higher_is_better"] is True
    assert j["value"] > 0 and j["steps"] == 1 and j["n_gpus"] == 1 and j["vs_baseline"] is None
    assert j["cpu_baseline"]["kind"] in ("port", "reference") and j["cpu_baseline"]["cores"] >= 1
    assert j["cpu_baseline"]["value"] == j["value"] and "sample" in j["cpu_baseline"]
    assert j["e2e"] == {"value": j["value"], "unit": "poses/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in j["config"] and j["metric"].startswith("poses/sec")


@pytest.mark.skipif(torch.cuda.is_available(), reason="needs a machine without a GPU")
def test_our_arm_has_no_cpu_fallback():
    p = run("--steps", "1", "--warmup", "0")
    assert p.returncode != 0 and "no CUDA device" in (p.stderr + p.stdout)
