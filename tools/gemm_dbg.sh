#!/bin/bash
for d in 0 1 2; do
PLR2_TC_DEBUG=$d ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -k regex:gemm_bf16x2 -c 6 --csv --log-file gpurun_out/dbg_$d.csv python bench.py --frames 64 --steps 1 --warmup 3 --profile-step > /dev/null 2>&1
done
