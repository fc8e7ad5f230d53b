#!/bin/bash
# Tuning builds with different NVCC flags: tools/build_flag_variants.sh name "flags" [name2 "flags2" ...]
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
