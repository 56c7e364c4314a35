#!/bin/bash
# Row-tile kernel (conv_rows.cuh): parity of the layers it takes over, the model-level tests, then the same-box A/B.
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-rows}
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "conv3x3_bn_every_shape or fused_head" > $OUT/${TAG}_parity.log 2>&1; echo "parity rc=$?"; tail -5 $OUT/${TAG}_parity.log
timeout 900 python -m pytest tests/test_gpu_model.py tests/test_gpu_frames.py -m gpu -q -x > $OUT/${TAG}_model.log 2>&1; echo "model rc=$?"; tail -5 $OUT/${TAG}_model.log
bash tools/gpu_variants.sh $TAG "ITMO_ROWS=0" "ITMO_ROWS_POOL=0"
