#!/bin/bash
# Scaling evidence on one multi-GPU box: bench.py at N = 1, 2, 4, 8 back to back (torchrun for N > 1).
# Usage (inside gpurun --gpus 8): tools/scale_round.sh <tag> [max_n]
tag=${1:-r1}; maxn=${2:-8}
mkdir -p gpurun_out
: > gpurun_out/scale_$tag.jsonl
for n in 1 2 4 8; do
  [ $n -gt $maxn ] && break
  if [ $n -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 >> gpurun_out/scale_$tag.jsonl 2> gpurun_out/scale_${tag}_n$n.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --steps 10 --warmup 3 >> gpurun_out/scale_$tag.jsonl 2> gpurun_out/scale_${tag}_n$n.err
  fi
  echo "N=$n rc=$?"
done
python - <<'P' "$tag"
import json, sys
rows = [json.loads(l) for l in open(f"gpurun_out/scale_{sys.argv[1]}.jsonl") if l.startswith("{")]
base = rows[0]["value"] if rows else None
for r in rows:
    print(r["n_gpus"], round(r["value"], 1), r["unit"], "x%.2f" % (r["value"] / base), "e2e", round(r["e2e"]["value"], 1))
P
