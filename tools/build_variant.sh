#!/bin/bash
# developer A/B builds: tools/build_variant.sh NAME [-DFLAG=...]...  ->  build/libwgpt_NAME.so  (use with WGPT_LIB=...)
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p build
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared "$@" \
  -o build/libwgpt_$name.so webgpu_path_tracer_b200/csrc/*.cu
cuobjdump -res-usage build/libwgpt_$name.so 2>/dev/null | grep -A1 "k_wf_extendILb0\|k_wf_shadeILb0" | grep -o "REG:[0-9]*\|STACK:[0-9]*\|k_wf_[a-z]*" | tr '\n' ' '; echo
