#!/bin/bash
# One GPU-box visit: parity tests, bench (default + sweeps), ncu launch list and one full capture.
# Usage (under gpurun): bash tools/gpu_round.sh [tag]
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $OUT/${TAG}_gpu.txt 2>&1
echo "== pytest -m gpu"; timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest_gpu.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $OUT/${TAG}_smoke.log
echo "== bench default"; timeout 600 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; cat $OUT/${TAG}_bench.json; tail -25 $OUT/${TAG}_bench.err
for c in 16 256; do
  echo "== bench chunk $c"; timeout 300 python bench.py --chunk $c --no-cpu --steps 3 > $OUT/${TAG}_bench_chunk$c.json 2> $OUT/${TAG}_bench_chunk$c.err; cat $OUT/${TAG}_bench_chunk$c.json | cut -c1-200
done
echo "== bench tf32"; timeout 300 python bench.py --precision tf32 --no-cpu --steps 3 --chunk 32 > $OUT/${TAG}_bench_tf32.json 2> $OUT/${TAG}_bench_tf32.err; cat $OUT/${TAG}_bench_tf32.json | cut -c1-300; tail -22 $OUT/${TAG}_bench_tf32.err
echo "== bench reference arm"; timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_ref.json 2>/dev/null; cat $OUT/${TAG}_bench_ref.json | cut -c1-200
SMALL="python bench.py --steps 1 --warmup 1 --batch 16 --chunk 16 --no-cpu"
echo "== ncu launches"
timeout 300 $SMALL > $OUT/${TAG}_small_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv $SMALL > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
echo "== ncu full (17 igemm launches of the timed step)"
timeout 300 $SMALL > $OUT/${TAG}_small_plain2.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:conv_igemm -s 51 -c 17 -f -o $OUT/${TAG}_igemm $SMALL > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $OUT
