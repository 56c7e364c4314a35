#!/bin/bash
# Timing experiments on ONE box: bash tools/gpu_env_sweep.sh <tag> "<ENV=..>" "<ENV=..>" ...   (bench args via BENCH_ARGS)
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-sw}; shift
i=0
for E in "$@"; do
  i=$((i+1))
  env $E timeout 600 python bench.py --no-cpu --steps 10 --batch 64 --chunk 64 $BENCH_ARGS > $OUT/${TAG}_$i.json 2> $OUT/${TAG}_$i.err
  echo "== $i [$E] $(python -c "import json;d=json.load(open('$OUT/${TAG}_$i.json'));print(d['value'],d['ms_per_step'],d['clocks']['sm_mhz'],d['clocks'].get('power_w_max'))" 2>&1 | tail -1)"
  grep -E "conv2d" $OUT/${TAG}_$i.err | awk '{printf "%s=%s ", $1, $2} END{print ""}'
done
