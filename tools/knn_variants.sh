#!/bin/bash
# Sweep of the compile-time knobs of kernel (4) on the GPU box (nvcc is in the image): rebuilds the library per
# variant and runs tools/knn_bench.py.  Usage (from the repo root): bash tools/knn_variants.sh > gpurun_out/knn_variants.log
set -u
run() {
  echo "=== $1"
  PLR2_EXTRA_NVCC_FLAGS="$1" python -m plr2_b200.build --force > /dev/null 2>&1 || { echo "build failed"; return; }
  PLR2_EXTRA_NVCC_FLAGS="$1" python tools/knn_bench.py 2>/dev/null | grep -v device | python -c "
import sys, json
for l in sys.stdin:
    j = json.loads(l); print('  %5d x %8d x %3d mode %d: %.2f Tpairs/s' % (j['n_ref'], j['n_query'], j['batch'], j['mode'], j['Tpairs_s']))"
}
run ""
run "-DPLR2_KNN_MINB=3"
run "-DPLR2_KNN_MINB=5"
run "-DPLR2_KNN_THREADS=256 -DPLR2_KNN_MINB=2"
run "-DPLR2_KNN_THREADS=64 -DPLR2_KNN_MINB=8"
run "-DPLR2_KNN_QPT=6 -DPLR2_KNN_MINB=5"
run "-DPLR2_KNN_QPT=4 -DPLR2_KNN_MINB=8"
run "-DPLR2_KNN_UNROLL=1"
run "-DPLR2_KNN_UNROLL=4"
python -m plr2_b200.build --force > /dev/null 2>&1
