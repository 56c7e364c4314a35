#!/bin/bash
# Training-step A/B on one box: bash tools/gpu_train_ab.sh <tag> "ENV=a" "ENV=b" ...   (first and last run = default)
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-tab}; shift
run() {
  name=$1; shift
  env "$@" timeout 600 python bench.py --workload train --steps 4 --warmup 2 > $OUT/${TAG}_$name.json 2> $OUT/${TAG}_$name.err
  python - <<PY
import json
l=[x for x in open('$OUT/${TAG}_$name.json').read().splitlines() if x.startswith('{')]
d=json.loads(l[-1]); ph=d['roofline']['phases']
print('== $name [$*]', d['value'], 'img/s', d['ms_per_step'], 'ms', d['clocks']['sm_mhz'], 'MHz |', {k:v['ms_per_step'] for k,v in ph.items()}, 'loss', d['train_loss'], 'frac', d['whole_step_frac_of_peak'])
PY
}
run base X=1
for v in "$@"; do run "$(echo $v | tr '= ' '__')" $v; done
run base2 X=1
