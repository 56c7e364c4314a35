#!/bin/bash
# Round-2 first visit: GPU tests, smoke, default bench (headline + secondary legs), reference arm.
TAG=${1:-r02a}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $OUT/${TAG}_gpu.txt 2>&1
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -E "Error|^FAILED|passed|failed" $OUT/${TAG}_pytest_gpu.log | tail -40
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -6 $OUT/${TAG}_smoke.log
echo "== bench default"; timeout 900 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; cat $OUT/${TAG}_bench.json; tail -60 $OUT/${TAG}_bench.err
