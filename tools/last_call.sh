#!/bin/bash
# tools/last_call.sh — one short gpurun call (1 GPU): the reference's ROOT- and TMinuit-using examples (built unmodified against tests/ref_root), the
# whole GPU test suite (no -x: every failure is listed), then the bench line and the reference arm. No ncu pass: no kernel changed
# since tools/final_capture.sh last ran.
set -u
R=$(pwd)
mkdir -p gpurun_out
T=$(mktemp -d)
for spec in "ref_dalitz --make-toy" "ref_addition" "ref_simpleFit" "ref_convolution" "ref_2d_plot" "ref_SigGen" "ref_TDDP4 -t 2" "ref_exponential1" "ref_exponential2" "ref_chisquare"; do
    set -- $spec
    name=$1; shift
    mkdir -p $T/$name
    s=$(date +%s.%N)
    (cd $T/$name && timeout 150 $R/examples/bin/$name "$@" > $R/gpurun_out/ex_$name.log 2>&1); rc=$?
    echo "example $name rc=$rc $(python -c "import time; print('%.1f s' % (time.time() - $s))")"
done
python tools/make_zachfit_toy.py $T/zach 100000
s=$(date +%s.%N)
(cd $T/zach && timeout 200 $R/examples/bin/ref_zachFit > $R/gpurun_out/ex_ref_zachFit.log 2>&1); echo "example ref_zachFit rc=$? $(python -c "import time; print('%.1f s' % (time.time() - $s))")"
timeout 420 python -m pytest tests -q -m gpu --durations=15 > gpurun_out/last_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/last_tests.log
timeout 300 python bench.py > gpurun_out/last_bench.json 2> gpurun_out/last_bench.err; echo "bench rc=$?"
timeout 120 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/last_ref.json 2> gpurun_out/last_ref.err; echo "ref rc=$?"
