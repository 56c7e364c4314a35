#!/bin/bash
# A/B timing of kernel variants on ONE box: bash tools/gpu_ab.sh <tag> "<ENV=a ...>" "<ENV=b ...>" [bench args]
# Runs bench.py (no CPU leg) alternately under the two environments, 2 rounds each, prints the per-layer tables.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-ab}; EA=$2; EB=$3; shift 3
for r in 1 2; do
  for v in A B; do
    if [ $v = A ]; then E=$EA; else E=$EB; fi
    env $E timeout 600 python bench.py --no-cpu --steps 10 --batch 64 --chunk 64 "$@" > $OUT/${TAG}_${v}${r}.json 2> $OUT/${TAG}_${v}${r}.err
    echo "== $v$r [$E] $(python -c "import json;d=json.load(open('$OUT/${TAG}_${v}${r}.json'));print(d['value'],d['ms_per_step'],d['clocks'])")"
    grep -E "conv2d|sum of" $OUT/${TAG}_${v}${r}.err | awk '{printf "%s=%s ", $1, $2} END{print ""}'
  done
done
