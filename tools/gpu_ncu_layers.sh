#!/bin/bash
# ncu --set full capture of one forward chunk (17 implicit-GEMM launches, 64 images) in the default precision, exported
# ON THE BOX to small CSVs (the .ncu-rep itself is too large for gpurun_out): the raw metric page of every launch and
# the SASS source page (stall sampling per instruction) of the launches named on the command line (default: conv2d_2,
# conv2d_transpose_4, conv2d_14 = launch indices 0, 15, 16).   usage: tools/gpu_ncu_layers.sh <tag> [idx ...]
set -u
tag=${1:-r02}
shift || true
idx=${@:-0 15 16}
rep=/tmp/${tag}_fwd
ncu --set full --import-source on --clock-control none -k regex:conv_igemm --launch-skip 51 --launch-count 17 -f -o $rep \
  python bench.py --batch 64 --steps 1 --warmup 1 --warm-seconds 0 --no-secondary --no-cpu --no-parity > gpurun_out/${tag}_ncu.log 2>&1
ncu -i $rep.ncu-rep --page raw --csv > gpurun_out/${tag}_fwd_raw.csv 2>> gpurun_out/${tag}_ncu.log
for i in $idx; do
  ncu -i $rep.ncu-rep --page source --csv --print-source sass --launch-skip $i --launch-count 1 > gpurun_out/${tag}_fwd_src_$i.csv 2>> gpurun_out/${tag}_ncu.log
  gzip -f gpurun_out/${tag}_fwd_src_$i.csv
done
ls -la gpurun_out/ | tail -8
