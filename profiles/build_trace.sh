#!/bin/sh
# Profiling build of the library with per-task timestamps in the replan-chain kernels (csrc/task_trace.cuh).
set -e
cd "$(dirname "$0")/.."
mkdir -p profiles/_trace
/usr/local/cuda/bin/nvcc -ccbin /usr/bin/g++ -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false \
  -Xcompiler -fPIC -shared -DHRG_TASK_TRACE -o profiles/_trace/libhrg_shield_trace.so human_robot_gym_b200/csrc/hrg_shield.cu
