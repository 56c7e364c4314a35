#!/bin/bash
# Training leg at several micro-batch sizes on one box: bash tools/gpu_train_mb.sh 32 64 32 64
for mb in "$@"; do
  timeout 400 python bench.py --workload train --micro-batch $mb --steps 3 --warmup 2 2>/dev/null > /tmp/mb_$mb.json
  python - <<PY
import json
d=json.loads([l for l in open('/tmp/mb_$mb.json') if l.startswith('{')][-1]); ph=d['roofline']['phases']
print('mb', $mb, d['value'], 'img/s', d['ms_per_step'], 'ms', d['clocks']['sm_mhz'], 'MHz', {k:v['ms_per_step'] for k,v in ph.items()}, 'e2e', d['e2e']['value'], 'loss', d['train_loss'])
PY
done
