#!/bin/bash
# Round evidence on one box: ncu launch list of a small default forward (+ secondary legs off), and the --set full
# capture of its 17 forward launches (fused stem + 16 implicit-GEMM) exported to the raw CSV page.   usage: tools/gpu_evidence.sh <tag>
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r02}
SMALL="python bench.py --batch 64 --steps 1 --warmup 1 --warm-seconds 0 --no-secondary --no-cpu --no-parity"
timeout 300 $SMALL > $OUT/${TAG}_small_plain.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/${TAG}_small_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv $SMALL > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"conv_igemm|rgb_conv|stem_fused|conv_rows" --launch-skip 51 --launch-count 17 -f -o /tmp/${TAG}_fwd $SMALL > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu -i /tmp/${TAG}_fwd.ncu-rep --page raw --csv > $OUT/${TAG}_fwd_raw.csv 2>> $OUT/${TAG}_ncu_full.log
ls -la $OUT | grep ${TAG}_
