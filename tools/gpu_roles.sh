#!/bin/bash
# Role ablation of conv_igemm_kernel on ONE box (results of the ablated runs are wrong by construction; timing only):
#   base | ITMO_TMA_STORE=0 | ITMO_DBG=32 (no epilogue work) | ITMO_DBG=64 (no MMAs) | ITMO_DBG=16 (no activation TMA after ring fill)
# usage: tools/gpu_roles.sh <tag> [extra "ENV=..." variants...]
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-roles}; shift
run() {
  name=$1; shift
  env "$@" timeout 300 python bench.py --no-cpu --no-secondary --no-parity --steps 8 --warm-seconds 1 > $OUT/${TAG}_$name.json 2> $OUT/${TAG}_$name.err
  echo "== $name [$*] $(python -c "import json;d=json.load(open('$OUT/${TAG}_$name.json'));print(d['value'],d['ms_per_step'],d['clocks']['sm_mhz'])")"
  grep -E "conv2d" $OUT/${TAG}_$name.err | awk '{printf "%s=%s ", $1, $2} END{print ""}'
}
run base X=1
run nostore ITMO_TMA_STORE=0
run noepi ITMO_DBG=32
run nomma ITMO_DBG=64
run noload ITMO_DBG=16
for v in "$@"; do run "$(echo $v | tr '= ' '__')" $v; done
run base2 X=1
