#!/bin/bash
# Same-box per-layer A/B of two builds: the tree's library against cnn-itmo_b200/lib_prev.so.alt (a copy of an earlier build).
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-lab}
L=cnn-itmo_b200/libitmo_b200.so
cp $L /tmp/lib_new.so
run() {
  name=$1
  timeout 300 python bench.py --no-cpu --no-secondary --no-parity --steps 8 --warm-seconds 1 > $OUT/${TAG}_$name.json 2> $OUT/${TAG}_$name.err
  echo "== $name $(python -c "import json;d=json.load(open('$OUT/${TAG}_$name.json'));print(d['value'],d['ms_per_step'],d['clocks']['sm_mhz'])")"
  grep -E "conv2d" $OUT/${TAG}_$name.err | awk '{printf "%s=%s ", $1, $2} END{print ""}'
}
run new1
cp cnn-itmo_b200/lib_prev.so.alt $L; touch $L
run prev1
cp /tmp/lib_new.so $L; touch $L
run new2
cp cnn-itmo_b200/lib_prev.so.alt $L; touch $L
run prev2
cp /tmp/lib_new.so $L; touch $L
