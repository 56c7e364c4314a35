#!/bin/bash
# Fused-exchange bring-up: single-GPU test, then the two-GPU bit-identity check and the training leg in both modes.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r02x}; N=${2:-2}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
timeout 300 python -m pytest tests/test_gpu_train.py -m gpu -q -x -k "fused_exchange or optimizer_update" 2>&1 | tail -4
timeout 600 $RUN tools/dp_check.py > $OUT/${TAG}_dp_check.txt 2>&1; echo "dp_check rc=$?"; tail -4 $OUT/${TAG}_dp_check.txt | cut -c1-1500
for m in fused nccl; do
  timeout 600 $RUN bench.py --gpus $N --workload train --steps 4 --warmup 2 --exchange $m > $OUT/${TAG}_train_$m.json 2> $OUT/${TAG}_train_$m.err; echo "train $m rc=$?"
  python - <<PY
import json
l=[x for x in open('$OUT/${TAG}_train_$m.json').read().splitlines() if x.startswith('{')]
d=json.loads(l[-1]); print('$m', d['value'], d['ms_per_step'], 'allreduce_ms', d['allreduce_ms'], d['exchange']['mode'], 'e2e', d['e2e']['value'], 'loss', d['train_loss'])
PY
done
