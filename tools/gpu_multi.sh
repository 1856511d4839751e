#!/usr/bin/env bash
# One multi-GPU call: TP parity tests for this box's GPU count, the bench line with its parity block and a
# per-call time line of one step, the collective microbenchmark. Everything lands in gpurun_out/.
#   gpurun --gpus N --timeout 1200 -- 'bash tools/gpu_multi.sh [pytest -k filter]'
set -u
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
mkdir -p gpurun_out
NG=$(nvidia-smi -L 2>/dev/null | wc -l)
FILTER="${1:-}"
python __graft_entry__.py > gpurun_out/build_n$NG.log 2>&1
if [ -n "$FILTER" ]; then
  FX_TP_TESTS=all timeout 900 python -m pytest tests/test_tp_gpu.py -x -q -m gpu -k "$FILTER" > gpurun_out/tests_tp_n$NG.log 2>&1
  echo "pytest rc=$?" >> gpurun_out/tests_tp_n$NG.log
fi
PORT=29531
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port $PORT \
  bench.py --gpus $NG --steps 20 --warmup 3 --trace gpurun_out/trace_n$NG.json > gpurun_out/bench_n$NG.json 2> gpurun_out/bench_n$NG.err
echo "bench rc=$?" >> gpurun_out/bench_n$NG.err
[ -n "${SKIP_SP:-}" ] || timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port $((PORT+1)) \
  bench.py --gpus $NG --steps 20 --warmup 3 --sequence-parallel 1 --trace gpurun_out/trace_n${NG}_sp.json > gpurun_out/bench_n${NG}_sp.json 2> gpurun_out/bench_n${NG}_sp.err
[ -n "${SKIP_SP:-}" ] || echo "bench rc=$?" >> gpurun_out/bench_n${NG}_sp.err
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port $((PORT+2)) \
  tools/bench_comm.py > gpurun_out/bench_comm_n$NG.log 2>&1
tail -3 gpurun_out/tests_tp_n$NG.log 2>/dev/null; tail -c 1500 gpurun_out/bench_n$NG.json; tail -2 gpurun_out/bench_n$NG.err; cat gpurun_out/bench_comm_n$NG.log | tail -6
