#!/bin/bash
# tools/final_capture.sh — everything profiles/ holds for a round, in one gpurun call (1 GPU): the GPU test suite, one ncu --set full
# capture of every hot kernel condensed into profiles/r02_ncu_full_summary.csv + profiles/traffic.json (so that the bench line that
# follows divides THIS build's flop and byte counts by its own CUDA-event times), the bench line, the reference arm and the ncu launch
# list of the headline bench. Every ncu pass runs only after the same command has exited 0 without it.
set -u
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/final_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/final_tests.log
python tools/profile_once.py > gpurun_out/final_profile_once.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none -k 'regex:k_amp|k_eval' -c 45 -f -o /tmp/final_full python tools/profile_once.py > gpurun_out/final_ncu_full.log 2>&1; echo "full capture rc=$?"
ncu -i /tmp/final_full.ncu-rep --page raw --csv > gpurun_out/final_full_raw.csv 2> gpurun_out/final_full_raw.err
python tools/ncu_summary.py gpurun_out/final_full_raw.csv profiles/r02_ncu_full_summary.csv profiles/traffic.json && cp profiles/r02_ncu_full_summary.csv profiles/traffic.json gpurun_out/
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/final_ref.json 2> gpurun_out/final_ref.err; echo "ref rc=$?"
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-secondary > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/final_launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/final_ncu_launches.log 2>&1; echo "launch list rc=$?"
ls -la gpurun_out | grep final_
