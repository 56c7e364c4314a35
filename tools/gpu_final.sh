#!/bin/bash
# Round-end visit on ONE GPU: parity tests, smoke, the three bench workloads + reference arm, ncu launch list of the
# default-size step's small twin and the --set full capture of one forward. bash tools/gpu_final.sh <tag>
TAG=${1:-r01_final}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $OUT/${TAG}_gpu.txt 2>&1
echo "== pytest -m gpu"; timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/${TAG}_pytest_gpu.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/${TAG}_smoke.log
echo "== bench reference arm"; timeout 400 python bench.py --impl reference > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err; echo "ref rc=$?"; cut -c1-300 $OUT/${TAG}_bench_ref.json
echo "== bench default"; timeout 600 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; cut -c1-400 $OUT/${TAG}_bench.json; grep -E "conv2d|sum of" $OUT/${TAG}_bench.err
echo "== bench 4k"; timeout 600 python bench.py --workload 4k > $OUT/${TAG}_bench_4k.json 2> $OUT/${TAG}_bench_4k.err; echo "4k rc=$?"; cut -c1-300 $OUT/${TAG}_bench_4k.json
echo "== bench train"; timeout 600 python bench.py --workload train > $OUT/${TAG}_bench_train.json 2> $OUT/${TAG}_bench_train.err; echo "train rc=$?"; cut -c1-300 $OUT/${TAG}_bench_train.json
echo "== bench tf32"; timeout 300 python bench.py --precision tf32 --no-cpu --steps 3 --chunk 32 > $OUT/${TAG}_bench_tf32.json 2> $OUT/${TAG}_bench_tf32.err; cut -c1-200 $OUT/${TAG}_bench_tf32.json
SMALL="python bench.py --steps 1 --warmup 3 --batch 64 --chunk 64 --no-cpu"
echo "== ncu launch list"
timeout 300 $SMALL > $OUT/${TAG}_small_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv $SMALL > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
echo "== ncu full (one forward: rgb conv + 17 implicit-GEMM launches)"
timeout 1200 ncu --set full --clock-control none -k regex:"conv_igemm|rgb_conv" -s 54 -c 18 -f -o $OUT/${TAG}_fwd $SMALL > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"; ls -la $OUT/${TAG}_fwd.ncu-rep
