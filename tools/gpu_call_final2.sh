#!/usr/bin/env bash
# Closing single-GPU call of round 2 on the final tree: the whole -m gpu suite, smoke, the bench
# lines kept in profiles/, and ncu captures of the 64-thread K1s builds that the automatic choice
# now uses at N = 256k.
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/final2
mkdir -p "$OUT"
PY=python
step() { local name=$1 secs=$2; shift 2; echo "== $name" | tee -a "$OUT/summary.txt"
	timeout "$secs" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "   exit $rc" | tee -a "$OUT/summary.txt"; return $rc; }
bench_line() { local name=$1 secs=$2; shift 2
	step "$name" "$secs" $PY bench.py "$@" && grep '^{' "$OUT/$name.log" | tail -1 > "$OUT/$name.json"; }
step pytest_gpu 1800 $PY -m pytest tests -m gpu -q -x --durations=10
step smoke 300 $PY __graft_entry__.py smoke
bench_line bench_default 600
bench_line bench_reference 600 --impl reference
bench_line bench_fp32 300 --precision 32 --steps 3 --warmup 3 --no-cpu-baseline
bench_line bench_256k 300 --bodies 262144 --steps 10 --warmup 3 --no-cpu-baseline
bench_line bench_20k 300 --bodies 20000 --dist uniform --steps 20 --warmup 5 --cpu-seconds 8
step ncu_launches_256k 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
	--log-file "$OUT/launches_256k.csv" $PY bench.py --bodies 262144 --steps 2 --warmup 1 --no-cpu-baseline --no-energy --no-parity --no-secondary
step ncu_full_sym64_v9_equal 600 ncu --set full --clock-control none --import-source on -k regex:force_sym_kernel -c 1 \
	-o "$OUT/k1s_v9_fp64_equal_262144" -f $PY tools/rhs_once.py --bodies 262144 --pair-symmetry 1 --reps 1
step ncu_full_sym64_v9_general 600 ncu --set full --clock-control none --import-source on -k regex:force_sym_kernel -c 1 \
	-o "$OUT/k1s_v9_fp64_general_262144" -f $PY tools/rhs_once.py --bodies 262144 --pair-symmetry 1 --unimass 0 --reps 1
step ncu_full_sym32_v5_general 600 ncu --set full --clock-control none --import-source on -k regex:force_sym_kernel -c 1 \
	-o "$OUT/k1s_v5_fp32_general_262144" -f $PY tools/rhs_once.py --precision 32 --bodies 262144 --pair-symmetry 1 --unimass 0 --reps 1
cat "$OUT/summary.txt"
