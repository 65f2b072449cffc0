set -u
R=$(pwd); mkdir -p gpurun_out; T=$(mktemp -d)
mkdir -p $T/a $T/b
(cd $T/a && timeout 200 $R/examples/bin/ref_SigGen > $R/gpurun_out/ex_ref_SigGen.log 2>&1); echo "example ref_SigGen rc=$?"
tail -3 gpurun_out/ex_ref_SigGen.log
(cd $T/b && timeout 200 $R/examples/bin/ref_TDDP4 -t 2 > $R/gpurun_out/ex_ref_TDDP4.log 2>&1); echo "example ref_TDDP4 rc=$?"
tail -5 gpurun_out/ex_ref_TDDP4.log
timeout 400 python -m pytest tests -q -m gpu --durations=8 > gpurun_out/last_tests2.log 2>&1; echo "tests rc=$?"; tail -14 gpurun_out/last_tests2.log
