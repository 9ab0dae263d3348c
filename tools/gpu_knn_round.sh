#!/bin/bash
# One GPU-box visit for kernel (4): KNN parity tests and the KNN micro-benchmark.
tag=${1:-x}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_knn_gpu.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/pytest_knn_$tag.log
timeout 300 python tools/knn_bench.py 2>&1 | tee gpurun_out/knn_bench_$tag.log
