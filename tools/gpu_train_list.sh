#!/bin/bash
# Training step launch list: ncu timing + DRAM bytes of every launch of ONE micro-batch of 16 images (serialised, side
# stream included).  usage: tools/gpu_train_list.sh <tag>
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r02train}
SMALL="python bench.py --workload train --train-batch 16 --micro-batch 16 --steps 1 --warmup 3 --exchange nccl"
timeout 300 $SMALL > $OUT/${TAG}_small_plain.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/${TAG}_small_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 640 -c 260 --csv --log-file $OUT/${TAG}_launches.csv $SMALL > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
