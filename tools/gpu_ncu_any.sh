#!/bin/bash
# ncu --set full capture of the launches matching a kernel-name regex in one small forward (64 images, default precision),
# exported ON THE BOX to small CSVs: raw metrics of every captured launch + the SASS source page of the launches named.
# usage: tools/gpu_ncu_any.sh <tag> <regex> <launch-skip> <launch-count> [idx ...]
set -u
tag=$1; rx=$2; skip=$3; cnt=$4; shift 4
idx=${@:-0}
rep=/tmp/${tag}
ncu --set full --import-source on --clock-control none -k regex:$rx --launch-skip $skip --launch-count $cnt -f -o $rep \
  python bench.py --batch 64 --steps 1 --warmup 1 --warm-seconds 0 --no-secondary --no-cpu --no-parity > gpurun_out/${tag}_ncu.log 2>&1
ncu -i $rep.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>> gpurun_out/${tag}_ncu.log
for i in $idx; do
  ncu -i $rep.ncu-rep --page source --csv --print-source sass --launch-skip $i --launch-count 1 > gpurun_out/${tag}_src_$i.csv 2>> gpurun_out/${tag}_ncu.log
  gzip -f gpurun_out/${tag}_src_$i.csv
done
ls -la gpurun_out/ | grep $tag
