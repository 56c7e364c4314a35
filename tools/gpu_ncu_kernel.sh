#!/bin/bash
# ncu --set full of selected launches of one small forward: bash tools/gpu_ncu_kernel.sh <tag> <kernel regex> <skip> <count>
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-k}; RE=${2:-rgb_conv}; SKIP=${3:-3}; CNT=${4:-1}
SMALL="python bench.py --steps 1 --warmup 3 --batch 64 --chunk 64 --no-cpu"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"$RE" -s $SKIP -c $CNT -f -o $OUT/${TAG} $SMALL > $OUT/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"; ls -la $OUT/${TAG}.ncu-rep
