#!/usr/bin/env bash
# Multi-GPU A/B of the tensor-parallel layer orders: FX_TP_ROWS / FX_TP_CHUNKS / FX_TP_ROWS_BLOCKS variants of bench.py
# (device-timed steps only), one JSON line each in gpurun_out/variants_n$NG.log, plus a time line of the default.
set -u
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}"
mkdir -p gpurun_out
NG=$(nvidia-smi -L 2>/dev/null | wc -l)
python __graft_entry__.py > gpurun_out/build_n$NG.log 2>&1
PORT=29600
run() {  # name, env...
  local name=$1; shift
  PORT=$((PORT+1))
  env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port $PORT \
    bench.py --gpus $NG --steps 20 --warmup 3 --profile-mode ${TRACE:+--trace gpurun_out/trace_n${NG}_$name.json} 2> gpurun_out/var_$name.err \
    | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(json.dumps({'variant':'$name','n':d['n_gpus'],'ms_per_step':round(d['ms_per_step'],4),'tokens_per_s':round(d['value'])}))" >> gpurun_out/variants_n$NG.log
}
: > gpurun_out/variants_n$NG.log
TRACE=1 run rows_c1 FX_TP_ROWS=1 FX_TP_CHUNKS=1
run replicated FX_TP_ROWS=0
run rows_c1_b16 FX_TP_ROWS=1 FX_TP_CHUNKS=1 FX_TP_ROWS_BLOCKS=16
run rows_c1_b64 FX_TP_ROWS=1 FX_TP_CHUNKS=1 FX_TP_ROWS_BLOCKS=64
cat gpurun_out/variants_n$NG.log
