#!/usr/bin/env bash
# One GPU call that re-establishes the state of the build on a fresh B200 and collects what the next
# optimisation decisions need. Everything lands in gpurun_out/ (merged back by gpurun).
#   gpurun --timeout 1500 -- 'bash tools/gpu_checkup.sh'
set -u
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
python __graft_entry__.py > gpurun_out/build.log 2>&1
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/tests_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/tests_gpu.log
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
# host-side launch cost: Python-sequenced kernels vs the native whole-layer entry vs a CUDA-graph replay
timeout 200 python tools/bench_native.py > gpurun_out/bench_native.log 2>&1
# candidate memory-bound kernels against the shipped ones (tools/micro/membound_v2.cu)
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/micro/membound_v2 tools/micro/membound_v2.cu -ldl \
  > gpurun_out/membound_build.log 2>&1 \
  && timeout 120 tools/micro/membound_v2 fleetx_b200/lib/libfleetx_b200.so 16384 2048 > gpurun_out/membound_v2.log 2>&1
timeout 200 python tools/bench_ops.py > gpurun_out/bench_ops.log 2>&1
tail -3 gpurun_out/tests_gpu.log; cat gpurun_out/bench_native.log gpurun_out/membound_v2.log 2>/dev/null | tail -20
