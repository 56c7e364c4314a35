#!/bin/bash
# Same-box A/B of the two builds (nvcc -split-compile 0 vs the single-threaded back end): training leg + inference headline.
# Needs cnn-itmo_b200/lib_nosplit.so.alt = the same sources built WITHOUT -split-compile 0 (nvcc flags of __graft_entry__.py).
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-split}
L=cnn-itmo_b200/libitmo_b200.so
cp $L /tmp/lib_split.so
run() {
  name=$1
  timeout 600 python bench.py --workload train --steps 4 --warmup 2 > $OUT/${TAG}_$name.json 2> $OUT/${TAG}_$name.err
  python - <<PY
import json
l=[x for x in open('$OUT/${TAG}_$name.json').read().splitlines() if x.startswith('{')]
d=json.loads(l[-1]); ph=d['roofline']['phases']
print('== $name', d['value'], 'img/s', d['ms_per_step'], 'ms', d['clocks']['sm_mhz'], 'MHz |', {k:v['ms_per_step'] for k,v in ph.items()}, 'loss', d['train_loss'])
PY
}
run split1
cp cnn-itmo_b200/lib_nosplit.so.alt $L; touch $L
run nosplit1
cp /tmp/lib_split.so $L; touch $L
run split2
cp cnn-itmo_b200/lib_nosplit.so.alt $L; touch $L
run nosplit2
timeout 600 python bench.py --no-secondary --no-cpu --steps 10 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('nosplit infer', d['value'], d['ms_per_step'], d['clocks']['sm_mhz'])"
cp /tmp/lib_split.so $L; touch $L
timeout 600 python bench.py --no-secondary --no-cpu --steps 10 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('split infer', d['value'], d['ms_per_step'], d['clocks']['sm_mhz'])"
