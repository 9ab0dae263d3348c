#!/bin/bash
# Evidence run for profiles/: bench line, launch list and full ncu capture of ONE default step (512 frames).
tag=${1:-r1}
mkdir -p gpurun_out
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err || { tail -5 gpurun_out/bench_$tag.err; exit 1; }
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$tag.json 2>> gpurun_out/bench_$tag.err
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$tag.csv \
  python bench.py --steps 1 --warmup 3 --profile-step > gpurun_out/ncu_l_$tag.log 2>&1
ncu --profile-from-start off --set full --clock-control none -o gpurun_out/step_$tag -f \
  python bench.py --steps 1 --warmup 3 --profile-step > gpurun_out/ncu_f_$tag.log 2>&1
tail -1 gpurun_out/ncu_f_$tag.log; ls -la gpurun_out/step_$tag.ncu-rep
# the report can exceed what gpurun copies back: keep the raw-metric CSV, drop the report if it is large
ncu -i gpurun_out/step_$tag.ncu-rep --page raw --csv > gpurun_out/step_${tag}_raw.csv 2>/dev/null
[ $(stat -c %s gpurun_out/step_$tag.ncu-rep) -gt 40000000 ] && rm -f gpurun_out/step_$tag.ncu-rep
du -sh gpurun_out
