#!/usr/bin/env bash
# After the cached step graph was keyed on what an upload can change: GPU suite + the small-N lines.
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/final3
mkdir -p "$OUT"
PY=python
step() { local name=$1 secs=$2; shift 2; echo "== $name" | tee -a "$OUT/summary.txt"
	timeout "$secs" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "   exit $rc" | tee -a "$OUT/summary.txt"; return $rc; }
bench_line() { local name=$1 secs=$2; shift 2
	step "$name" "$secs" $PY bench.py "$@" && grep '^{' "$OUT/$name.log" | tail -1 > "$OUT/$name.json"; }
step pytest_gpu 1800 $PY -m pytest tests -m gpu -q -x
step smoke 300 $PY __graft_entry__.py smoke
bench_line bench_20k 300 --bodies 20000 --dist uniform --steps 20 --warmup 5 --cpu-seconds 8
bench_line bench_64k 300 --bodies 65536 --steps 20 --warmup 5 --no-cpu-baseline
bench_line bench_default 600
step latency_4body 120 $PY tools/latency_4body.py
cat "$OUT/summary.txt"
