#!/bin/bash
# ncu visit: launch list + full capture of the 17 igemm launches of one small forward. bash tools/gpu_ncu.sh <tag>
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-ncu}
SMALL="python bench.py --steps 1 --warmup 1 --batch 16 --chunk 16 --no-cpu"
timeout 300 $SMALL > $OUT/${TAG}_small_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv $SMALL > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 300 $SMALL > $OUT/${TAG}_small_plain2.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"conv_igemm|rgb" -s 54 -c 18 -f -o $OUT/${TAG}_fwd $SMALL > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
