#!/bin/bash
# Per-layer timing of bench.py's headline leg under several environments on ONE box (first and last run = default).
# usage: tools/gpu_variants.sh <tag> "ENV=a" "ENV=b ENV2=c" ...
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-var}; shift
run() {
  name=$1; shift
  env "$@" timeout 300 python bench.py --no-cpu --no-secondary --no-parity --steps 8 --warm-seconds 1 > $OUT/${TAG}_$name.json 2> $OUT/${TAG}_$name.err
  echo "== $name [$*] $(python -c "import json;d=json.load(open('$OUT/${TAG}_$name.json'));print(d['value'],d['ms_per_step'],d['clocks']['sm_mhz'])")"
  grep -E "conv2d" $OUT/${TAG}_$name.err | awk '{printf "%s=%s ", $1, $2} END{print ""}'
}
run base X=1
for v in "$@"; do run "$(echo $v | tr '= ' '__')" $v; done
run base2 X=1
