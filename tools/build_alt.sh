#!/bin/bash
# tools/build_alt.sh NAME "-DMACRO=.. ..." — an alternative build of libgoofit_b200.so with extra nvcc flags, into
# alt_builds/NAME/libgoofit_b200.so, for tools/ab_compare.sh (timed back to back with the in-tree library on one box)
set -e
name=$1; shift
out=alt_builds/$name
mkdir -p $out
for f in goofit_b200/csrc/*.cu; do
  b=$(basename $f .cu)
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr "$@" -c $f -o $out/$b.o &
done
wait
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $out/libgoofit_b200.so $out/*.o -lcudart -ldl
echo built $out/libgoofit_b200.so
