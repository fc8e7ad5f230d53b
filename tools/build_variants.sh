#!/bin/bash
# Tuning builds of libgrt.so: tools/build_variants.sh name "-DGRT_X=.. -DGRT_Y=.." [name2 "flags2" ...]
# -> gaussian-ray-tracing_b200/csrc/variants/libgrt_<name>.so ; select with GRT_LIBGRT=<path>.
set -e
cd "$(dirname "$0")/../gaussian-ray-tracing_b200/csrc"
mkdir -p variants
ARCH="-gencode arch=compute_100a,code=sm_100a"
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  (nvcc $ARCH -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $flags -c trace.cu -o variants/trace_$name.o && \
   nvcc $ARCH -shared -o variants/libgrt_$name.so capi.o build.o ingest.o variants/trace_$name.o -ldl && rm variants/trace_$name.o && echo built $name) &
done
wait
