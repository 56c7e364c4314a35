#!/bin/bash
# Quick GPU visit: bash tools/gpu_quick.sh <tag> "<pytest -k expr>" [bench args...]
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-q}; KEXPR=${2:-}; shift 2
if [ -n "$KEXPR" ]; then
  timeout 1500 python -m pytest tests -m gpu -q -k "$KEXPR" > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"
else
  timeout 1500 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"
fi
grep -E "AssertionError|^FAILED|passed|failed" $OUT/${TAG}_pytest.log | tail -40
timeout 600 python bench.py --no-cpu --steps 3 "$@" > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; cut -c1-300 $OUT/${TAG}_bench.json; tail -22 $OUT/${TAG}_bench.err
