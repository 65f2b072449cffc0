#!/bin/bash
# tools/profile_source.sh <kernel regex> <launch skip> <launch count> <profile_once workload> [tag]
# One ncu --set full capture with source correlation of the selected launches of tools/profile_once.py at 10^7 events, exported
# ON the box (the .ncu-rep is too big for gpurun_out/) as a per-SASS-instruction CSV and the raw metric page.
# Read with tools/sass_by_line.py against `nvdisasm -g -c` of the cubin (cuobjdump -xelf all build/<file>.o).
set -u
rx=$1; skip=$2; cnt=$3; kind=$4; tag=${5:-src}
mkdir -p gpurun_out
timeout 900 ncu --set full --import-source on --clock-control none -k regex:$rx --launch-skip $skip -c $cnt -f -o /tmp/$tag \
    python tools/profile_once.py 10000000 $kind > gpurun_out/$tag.log 2>&1
echo "capture rc=$?"
ncu -i /tmp/$tag.ncu-rep --page source --print-source sass --csv > gpurun_out/${tag}_sass.csv 2> gpurun_out/${tag}_sass.err
ncu -i /tmp/$tag.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2> gpurun_out/${tag}_raw.err
ls -la gpurun_out | grep $tag
