#!/usr/bin/env bash
# Final single-GPU call of round 2: whole -m gpu suite on the final build, the bench lines that go
# into profiles/, ncu evidence at the headline size, the MUL+ADD accumulate experiment.
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/final1
mkdir -p "$OUT"
PY=python
step() { local name=$1 secs=$2; shift 2; echo "== $name" | tee -a "$OUT/summary.txt"
	timeout "$secs" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "   exit $rc" | tee -a "$OUT/summary.txt"; return $rc; }
bench_line() { local name=$1 secs=$2; shift 2
	step "$name" "$secs" $PY bench.py "$@" && grep '^{' "$OUT/$name.log" | tail -1 > "$OUT/$name.json"; }
step pytest_gpu 1800 $PY -m pytest tests -m gpu -q -x --durations=15
step smoke 300 $PY __graft_entry__.py smoke
bench_line bench_default 600
bench_line bench_reference 600 --impl reference
for v in 9 10 11; do
	bench_line bench_acc_v$v 300 --sym-variant $v --steps 2 --warmup 3 --no-cpu-baseline --no-energy --no-secondary
done
bench_line bench_fp32 300 --precision 32 --steps 3 --warmup 3 --no-cpu-baseline
bench_line bench_256k 300 --bodies 262144 --steps 10 --warmup 3 --no-cpu-baseline
bench_line bench_20k 300 --bodies 20000 --dist uniform --steps 20 --warmup 5 --cpu-seconds 8
bench_line bench_20k_k1 300 --bodies 20000 --dist uniform --pair-symmetry 0 --steps 20 --warmup 5 --no-cpu-baseline
bench_line bench_k1 400 --pair-symmetry 0 --steps 2 --warmup 3 --no-cpu-baseline
step latency_4body 120 $PY tools/latency_4body.py
step ncu_launches_1m 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
	--log-file "$OUT/launches_1m.csv" $PY bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-energy --no-parity --no-secondary
step ncu_launches_20k 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
	--log-file "$OUT/launches_20k.csv" $PY bench.py --bodies 20000 --dist uniform --steps 3 --warmup 1 --no-cpu-baseline --no-energy --no-parity --no-secondary
step ncu_full_sym64_1m 900 ncu --set full --clock-control none --import-source on -k regex:force_sym_kernel -c 1 \
	-o "$OUT/k1s_fp64_equal_1m" -f $PY tools/rhs_once.py --bodies 1048576 --reps 1
step ncu_full_sym64_20k 300 ncu --set full --clock-control none --import-source on -k regex:force_sym_kernel -c 1 \
	-o "$OUT/k1s_fp64_general_20000" -f $PY tools/rhs_once.py --bodies 20000 --unimass 0
cat "$OUT/summary.txt"
