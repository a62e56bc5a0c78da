#!/usr/bin/env bash
# Third GPU call of round 2 (one B200): K1s with strip-granular slices at small and medium N.
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/call3
mkdir -p "$OUT"
PY=python
step() { local name=$1 secs=$2; shift 2; echo "== $name" | tee -a "$OUT/summary.txt"
	timeout "$secs" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "   exit $rc" | tee -a "$OUT/summary.txt"; return $rc; }
bench_line() { local name=$1 secs=$2; shift 2
	step "$name" "$secs" $PY bench.py "$@" && grep '^{' "$OUT/$name.log" | tail -1 > "$OUT/$name.json"; }
step pytest_sym 900 $PY -m pytest tests/test_zzz_gpu_pair_symmetry.py tests/test_zzzz_gpu_bounds_check.py tests/test_zzzz_gpu_headline_size.py -m gpu -q -x
step crossover 900 $PY tools/sym_crossover.py --sizes 4096,8192,12288,16384,20000,24576,32768,40960,49152,65536,98304 \
	--variants64 -1,8,7,5,3 --variants32 -1,1,3,2
grep '^{' "$OUT/crossover.log" > "$OUT/crossover.jsonl"
bench_line bench_20k_sym 300 --bodies 20000 --dist uniform --pair-symmetry 1 --steps 20 --warmup 5 --no-cpu-baseline
bench_line bench_20k_sym_v5 300 --bodies 20000 --dist uniform --pair-symmetry 1 --sym-variant 5 --steps 20 --warmup 5 --no-cpu-baseline
bench_line bench_default 400 --steps 3 --warmup 3 --no-cpu-baseline --trace-out "$OUT/trace_1gpu_1m.json"
bench_line bench_256k 300 --bodies 262144 --steps 5 --warmup 3 --no-cpu-baseline
cat "$OUT/summary.txt"
