#!/bin/bash
# tools/run_ref_python_examples.sh stage|run — the reference's own python/examples scripts, UNMODIFIED, against this repository's
# `goofit` package on the GPU. `stage` (here, where /root/reference exists) copies the scripts into examples/bin/pyref, a git-ignored
# scratch directory that travels to the GPU box with the built binaries and is deleted again after the call: nothing of the
# reference enters the repository. `run` (on the box) executes each script with this repository and tests/ref_mpl (a headless
# matplotlib stand-in) on PYTHONPATH and prints its exit status. Left out: DP4.py and dalitz.py (they read data files of the
# reference checkout's build tree).
set -u
R=$(cd "$(dirname "$0")/.." && pwd)
S=$R/examples/bin/pyref
if [ "$1" = stage ]; then
    rm -rf $S; mkdir -p $S
    for f in exponential product convolution addition chisquare plot_2d simpleFit SigGen TDDP4 generate_sig_example quad1f zachFit; do
        cp /root/reference/python/examples/$f.py $S/
    done
    ls $S | tr '\n' ' '; echo
    exit 0
fi
mkdir -p $R/gpurun_out
W=$(mktemp -d)
python $R/tools/make_zachfit_toy.py $W/examples/zachFit 100000
cd $W
for f in exponential product convolution addition chisquare plot_2d simpleFit SigGen TDDP4 generate_sig_example quad1f zachFit; do
    s=$(date +%s)
    PYTHONPATH=$R:$R/tests/ref_mpl timeout 100 python $S/$f.py > $R/gpurun_out/py_$f.log 2>&1; rc=$?
    echo "python example $f.py rc=$rc $(( $(date +%s) - s )) s"
    [ $rc -ne 0 ] && tail -4 $R/gpurun_out/py_$f.log | cut -c1-300
done
exit 0
