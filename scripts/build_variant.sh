#!/bin/bash
# dev aid: build libngp_b200 variants with extra -D flags:  [VFILES="render.cu"] scripts/build_variant.sh NAME -DFQ_HINT=1 ...
# (only the files in $VFILES, default field.cu, are recompiled with the flags)
set -e
name=$1; shift
cd "$(dirname "$0")/../nerf_navigation_b200"
mkdir -p build/var_$name build/variants
for f in csrc/*.cu; do
  o=build/var_$name/$(basename ${f%.cu}).o
  if [[ " ${VFILES:-field.cu} " == *" $(basename $f) "* ]] || [ ! -f build/$(basename ${f%.cu}).o ]; then
    nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr "$@" -c $f -o $o &
  else
    cp build/$(basename ${f%.cu}).o $o
  fi
done
wait
nvcc -shared -o build/variants/libngp_$name.so build/var_$name/*.o -gencode arch=compute_100a,code=sm_100a
echo built nerf_navigation_b200/build/variants/libngp_$name.so
