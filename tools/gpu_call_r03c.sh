#!/usr/bin/env bash
# Round-3 call C (G GPUs, default 1): the default bench line with its new blocks (parity after one
# Euler step, secondary.configs), the reference arm, and on 2 GPUs the odd-block configuration.
set -u
cd "$(dirname "$0")/.."
G=${1:-1}
OUT=gpurun_out/r03c_$G
mkdir -p "$OUT"
PY=python
step() { local name=$1 secs=$2; shift 2; echo "== $name" | tee -a "$OUT/summary.txt"
	timeout "$secs" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "   exit $rc" | tee -a "$OUT/summary.txt"; return $rc; }
if [ "$G" = 1 ]; then
	step bench_default 600 $PY bench.py && grep '^{' "$OUT/bench_default.log" | tail -1 > "$OUT/bench_default.json"
	step bench_reference 400 $PY bench.py --impl reference && grep '^{' "$OUT/bench_reference.log" | tail -1 > "$OUT/bench_reference.json"
	step bench_fp32 400 $PY bench.py --precision 32 --steps 3 --warmup 3 --no-cpu-baseline && grep '^{' "$OUT/bench_fp32.log" | tail -1 > "$OUT/bench_fp32.json"
	step pytest_gpu 900 $PY -m pytest tests -m gpu -x -q
else
	TR="$PY -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1"
	step bench_default 600 $TR --master-port 29571 bench.py --gpus $G --steps 3 --warmup 3 && grep '^{' "$OUT/bench_default.log" | tail -1 > "$OUT/bench_default.json"
	step bench_1100000_general 400 $TR --master-port 29572 bench.py --gpus $G --bodies 1100000 --steps 2 --warmup 3 --unimass 0 --no-secondary && grep '^{' "$OUT/bench_1100000_general.log" | tail -1 > "$OUT/bench_1100000_general.json"
fi
cat "$OUT/summary.txt"
