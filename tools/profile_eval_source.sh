#!/bin/bash
# tools/profile_eval_source.sh <dp4|tddp4|dalitz> — one ncu --set full capture of k_eval_sum on a 10^7-event workload with
# source correlation, exported ON the box (the .ncu-rep is too big for gpurun_out/) as per-CUDA-line and per-SASS CSVs.
set -u
kind=${1:-dp4}
mkdir -p gpurun_out
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_eval_sum -c 2 -f -o /tmp/ev_$kind \
    python tools/profile_once.py 10000000 $kind > gpurun_out/ev_$kind.log 2>&1
echo "capture rc=$?"
ncu -i /tmp/ev_$kind.ncu-rep --page source --print-source cuda --csv > gpurun_out/ev_${kind}_cuda.csv 2> gpurun_out/ev_${kind}_cuda.err
ncu -i /tmp/ev_$kind.ncu-rep --page source --print-source sass --csv > gpurun_out/ev_${kind}_sass.csv 2> gpurun_out/ev_${kind}_sass.err
ls -la gpurun_out | grep ev_
