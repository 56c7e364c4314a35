#!/bin/bash
# Training step: ncu launch list (one micro-batch of 16 images) and a full capture of the wgrad + a few elementwise launches.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r01_train}
SMALL="python bench.py --workload train --train-batch 16 --micro-batch 16 --steps 1 --warmup 3"
timeout 300 $SMALL > $OUT/${TAG}_small_plain.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/${TAG}_small_plain.log; exit 1; }
# step = ~193 launches; skip the setup (weight sync ~40) and the 3 warm-up steps, list one whole timed step
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 640 -c 230 --csv --log-file $OUT/${TAG}_launches.csv $SMALL > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 300 $SMALL > $OUT/${TAG}_small_plain2.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none -k regex:wgrad_igemm -s 78 -c 26 -f -o $OUT/${TAG}_wgrad $SMALL > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $OUT | tail -8
