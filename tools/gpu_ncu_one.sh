#!/bin/bash
# Full ncu capture of ONE igemm launch of a small forward: bash tools/gpu_ncu_one.sh <tag> <launch index 0..16 within the forward>
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-one}; IDX=${2:-16}
SMALL="python bench.py --steps 1 --warmup 1 --batch 16 --chunk 16 --no-cpu"
timeout 300 $SMALL > $OUT/${TAG}_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_igemm -s $((51 + IDX)) -c 1 -f -o $OUT/${TAG} $SMALL > $OUT/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"
