#!/bin/bash
# ncu --set full of the 17 implicit-GEMM launches of ONE forward at the bench's arena chunk (64 images): DRAM traffic per launch
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r01_traffic}
SMALL="python bench.py --steps 1 --warmup 3 --batch 64 --chunk 64 --no-cpu"
timeout 300 $SMALL > $OUT/${TAG}_plain.log 2>&1 || { echo plain failed; exit 1; }
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"conv_igemm|rgb_conv" -s 54 -c 18 -f -o $OUT/${TAG}_igemm $SMALL > $OUT/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"; ls -la $OUT/${TAG}_igemm.ncu-rep
