#!/bin/bash
# The driver's launch line for the default bench at N ranks (under gpurun --gpus N): bash tools/gpu_nbench.sh <tag> <N>
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-nb}; N=${2:-8}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $RUN bench.py --gpus $N --steps 20 --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench n$N rc=$?"
python - <<PY
import json
l=[x for x in open('$OUT/${TAG}_bench.json').read().splitlines() if x.startswith('{')]
d=json.loads(l[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'parity',d['parity']['psnr_db'],'roofline',d['roofline']['frac'],d['clocks'])
for k,v in d.get('secondary',{}).items(): print(k, {kk:v.get(kk) for kk in ('value','ms_per_step','allreduce_ms','h2d_ms','error')}, 'e2e', (v.get('e2e') or {}))
PY
tail -3 $OUT/${TAG}_bench.err
