#!/bin/bash
# CTA-pair bring-up: parity tests under a timeout, then per-layer A/B against ITMO_PAIR=0.
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -q -x -k "parity or model" > $OUT/pair_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/pair_pytest.log | cut -c1-300
bash tools/gpu_variants.sh r02pair ITMO_PAIR=0
