#!/bin/bash
# N-GPU training visit: DP consistency check + bench.py --workload train at N (and the inference headline at N).
N=${1:-2}; OUT=gpurun_out; mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $RUN tools/dp_check.py > $OUT/dp_check_n$N.log 2>&1; echo "dp_check rc=$?"; grep -E "^\[|^\{|dp_check" $OUT/dp_check_n$N.log | tail -3
timeout 900 $RUN bench.py --gpus $N --workload train --steps 5 --warmup 3 > $OUT/train_n$N.json 2> $OUT/train_n$N.err; echo "bench train rc=$?"; cut -c1-600 $OUT/train_n$N.json; tail -3 $OUT/train_n$N.err
timeout 600 python bench.py --workload train --steps 5 --warmup 3 > $OUT/train_n1.json 2> $OUT/train_n1.err; echo "bench train n1 rc=$?"; cut -c1-400 $OUT/train_n1.json
