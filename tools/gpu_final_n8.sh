#!/usr/bin/env bash
# The 8-GPU evidence run: TP parity tests at world 8, the bench line (with its parity block and a time line) for the
# three BASELINE layer shapes at mp 8. Everything lands in gpurun_out/.
set -u
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
mkdir -p gpurun_out
NG=$(nvidia-smi -L 2>/dev/null | wc -l)
python __graft_entry__.py > gpurun_out/build_n$NG.log 2>&1
FX_TP_TESTS=all timeout 900 python -m pytest tests/test_tp_gpu.py -x -q -m gpu -k "${1:-8-}" > gpurun_out/tests_tp_n$NG.log 2>&1
echo "pytest rc=$?" >> gpurun_out/tests_tp_n$NG.log
PORT=29700
for cfg in c3 c4 c5; do
  PORT=$((PORT+1))
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port $PORT \
    bench.py --gpus $NG --steps 20 --warmup 3 --config $cfg --trace gpurun_out/trace_n${NG}_$cfg.json \
    > gpurun_out/bench_n${NG}_$cfg.json 2> gpurun_out/bench_n${NG}_$cfg.err
  echo "bench $cfg rc=$?" >> gpurun_out/bench_n${NG}_$cfg.err
done
tail -3 gpurun_out/tests_tp_n$NG.log
for cfg in c3 c4 c5; do tail -1 gpurun_out/bench_n${NG}_$cfg.err; tail -c 600 gpurun_out/bench_n${NG}_$cfg.json; echo; done
