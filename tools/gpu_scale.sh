#!/bin/bash
# Scaling visit at N GPUs (under gpurun --gpus N): the three bench workloads + the reference arm, launched as the driver does.
N=${1:-4}; OUT=gpurun_out; mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
timeout 600 $RUN bench.py --gpus $N --steps 10 --warmup 3 > $OUT/scale_batch_n$N.json 2> $OUT/scale_batch_n$N.err; echo "batch n$N rc=$?"; cut -c1-260 $OUT/scale_batch_n$N.json
timeout 600 $RUN bench.py --gpus $N --workload 4k --steps 3 > $OUT/scale_4k_n$N.json 2> $OUT/scale_4k_n$N.err; echo "4k n$N rc=$?"; cut -c1-260 $OUT/scale_4k_n$N.json
timeout 900 $RUN bench.py --gpus $N --workload train --steps 5 --warmup 3 > $OUT/scale_train_n$N.json 2> $OUT/scale_train_n$N.err; echo "train n$N rc=$?"; cut -c1-260 $OUT/scale_train_n$N.json
timeout 300 $RUN bench.py --gpus $N --impl reference --steps 2 --warmup 1 > $OUT/scale_ref_n$N.json 2> $OUT/scale_ref_n$N.err; echo "ref n$N rc=$?"; cut -c1-200 $OUT/scale_ref_n$N.json
tail -2 $OUT/scale_*_n$N.err | tail -16
