#!/usr/bin/env bash
# Round-3 two-GPU call: the multi-GPU GPU tests (skipped on a one-GPU box), the parity check, and
# the configuration the round-2 advisor named: a per-rank block that is an ODD multiple of 256 rows
# at a size where the preferred builds have 512-/1024-row blocks (n = 1 100 000 on 2 GPUs).
#   gpurun --gpus 2 --timeout 1200 -- 'bash tools/gpu_call_r03_multi.sh'
set -u
cd "$(dirname "$0")/.."
G=2
OUT=gpurun_out/r03m
mkdir -p "$OUT"
PY=python
step() { local name=$1 secs=$2; shift 2; echo "== $name" | tee -a "$OUT/summary.txt"
	timeout "$secs" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "   exit $rc" | tee -a "$OUT/summary.txt"; return $rc; }
TR="$PY -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1"
bench_line() { local name=$1 port=$2 secs=$3; shift 3
	step "$name" "$secs" $TR --master-port $port bench.py --gpus $G "$@" && grep '^{' "$OUT/$name.log" | tail -1 > "$OUT/$name.json"; }
step pytest_gpu_multi 600 $PY -m pytest tests -m gpu -x -q
bench_line bench_1100000 29561 400 --bodies 1100000 --steps 2 --warmup 3 --no-secondary
bench_line bench_1100000_det 29562 400 --bodies 1100000 --steps 2 --warmup 3 --deterministic 1 --no-secondary
bench_line bench_1100000_general 29563 400 --bodies 1100000 --steps 2 --warmup 3 --unimass 0 --no-secondary
bench_line bench_1m 29564 400 --steps 3 --warmup 3 --no-secondary
NBODY_PAIR_SYM=auto step multi_check_sym_auto 600 $TR --master-port 29541 tests/tools/multi_check.py 6000
cat "$OUT/summary.txt"
