#!/bin/bash
# ncu --set full of the elementwise kernels of ONE training micro-batch (16 images): bash tools/gpu_train_ncu_elem.sh <tag>
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r01_train_elem}
SMALL="python bench.py --workload train --train-batch 16 --micro-batch 16 --steps 1 --warmup 3"
timeout 1500 ncu --section SpeedOfLight --section MemoryWorkloadAnalysis --section WarpStateStats --section Occupancy --section LaunchStats --section SchedulerStats --clock-control none -k regex:"bn_bwd_apply|bn_bwd_reduce|bn_apply|head_train|pool_bwd" -s 192 -c 64 -f -o $OUT/${TAG} $SMALL > $OUT/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"; ls -la $OUT/${TAG}.ncu-rep
