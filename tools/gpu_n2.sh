#!/bin/bash
# Two-GPU visit (under gpurun --gpus 2): the driver's own launch line for the default bench at N = 2, the reference arm,
# the two-GPU data-parallel bit-identity check, and the GPU frames tests.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r02n2}; N=${2:-2}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $RUN bench.py --gpus $N --steps 20 --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench n$N rc=$?"
python - <<PY
import json
d=json.load(open('$OUT/${TAG}_bench.json'))
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'parity',d['parity']['psnr_db'],'roofline',d['roofline']['frac'],d['clocks'])
for k,v in d.get('secondary',{}).items(): print(k, {kk:v.get(kk) for kk in ('value','ms_per_step','allreduce_ms','h2d_ms','error')}, 'e2e', (v.get('e2e') or {}))
PY
timeout 300 $RUN bench.py --gpus $N --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_ref.json 2> $OUT/${TAG}_ref.err; echo "ref n$N rc=$?"; cut -c1-300 $OUT/${TAG}_ref.json
timeout 600 $RUN tools/dp_check.py > $OUT/${TAG}_dp_check.txt 2>&1; echo "dp_check rc=$?"; tail -8 $OUT/${TAG}_dp_check.txt
timeout 600 python -m pytest tests -m gpu -q -x -k "frames" 2>&1 | tail -3
