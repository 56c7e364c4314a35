#!/bin/bash
# Multi-GPU visit: bash tools/gpu_multi.sh <tag> <ngpus>   (under gpurun --gpus N)
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-mg}; N=${2:-2}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 python bench.py --workload 4k --steps 3 > $OUT/${TAG}_4k_n1.json 2> $OUT/${TAG}_4k_n1.err; echo "4k n1 rc=$?"; cut -c1-330 $OUT/${TAG}_4k_n1.json
timeout 600 $RUN bench.py --gpus $N --workload 4k --steps 3 > $OUT/${TAG}_4k_n$N.json 2> $OUT/${TAG}_4k_n$N.err; echo "4k n$N rc=$?"; cut -c1-330 $OUT/${TAG}_4k_n$N.json
timeout 600 python bench.py --no-cpu --steps 10 > $OUT/${TAG}_batch_n1.json 2> $OUT/${TAG}_batch_n1.err; echo "batch n1 rc=$?"; cut -c1-330 $OUT/${TAG}_batch_n1.json
timeout 600 $RUN bench.py --gpus $N --steps 10 > $OUT/${TAG}_batch_n$N.json 2> $OUT/${TAG}_batch_n$N.err; echo "batch n$N rc=$?"; cut -c1-330 $OUT/${TAG}_batch_n$N.json
timeout 300 $RUN bench.py --gpus $N --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_ref_n$N.json 2> $OUT/${TAG}_ref_n$N.err; echo "ref n$N rc=$?"; cut -c1-200 $OUT/${TAG}_ref_n$N.json
tail -3 $OUT/${TAG}_*.err | tail -30
