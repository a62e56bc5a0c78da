#!/usr/bin/env bash
# Second GPU call of round 2 (one B200): whole -m gpu suite on the final build, the headline bench
# line, crossover sweep, ncu evidence (K1s default build, N = 20k), reference GPU kernel rows.
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/call2
mkdir -p "$OUT"
PY=python
step() { local name=$1 secs=$2; shift 2; echo "== $name" | tee -a "$OUT/summary.txt"
	timeout "$secs" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "   exit $rc" | tee -a "$OUT/summary.txt"; return $rc; }
bench_line() { local name=$1 secs=$2; shift 2
	step "$name" "$secs" $PY bench.py "$@" && grep '^{' "$OUT/$name.log" | tail -1 > "$OUT/$name.json"; }

step pytest_gpu 1500 $PY -m pytest tests -m gpu -q -x
step pytest_gpu_rest 1500 $PY -m pytest tests -m gpu -q --deselect tests/test_gpu_parity.py
bench_line bench_default 600 --steps 5 --warmup 3
bench_line bench_k1 400 --pair-symmetry 0 --steps 2 --warmup 3 --no-cpu-baseline --no-secondary
bench_line bench_fp32 300 --precision 32 --steps 3 --warmup 3 --no-cpu-baseline
bench_line bench_256k 300 --bodies 262144 --steps 5 --warmup 3 --no-cpu-baseline
bench_line bench_20k 300 --bodies 20000 --dist uniform --steps 20 --warmup 5 --cpu-seconds 8
step crossover 900 $PY tools/sym_crossover.py
grep '^{' "$OUT/crossover.log" > "$OUT/crossover.jsonl"
# reference GPU kernel (lab_5 N_body.cu recompiled for sm_100), FP32 as committed and FP64 typedefs
step ref_gpu_fp32 300 oracle/_ref/lab5_ref 262144 2
step ref_gpu_fp64 600 oracle/_ref/lab5_ref_fp64 262144 1
step ref_gpu_fp32_1m 300 oracle/_ref/lab5_ref 1048576 1
# ncu: launch lists, then full captures (each only after its command ran clean above)
step ncu_launches_1m 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
	--log-file "$OUT/launches_1m.csv" $PY bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-energy --no-parity --no-secondary
step ncu_launches_20k 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
	--log-file "$OUT/launches_20k.csv" $PY bench.py --bodies 20000 --dist uniform --steps 3 --warmup 1 --no-cpu-baseline --no-energy --no-parity --no-secondary
step ncu_full_sym64 600 ncu --set full --clock-control none --import-source on -k regex:force_sym_kernel -c 1 \
	-o "$OUT/k1s_fp64_equal_262144" -f $PY tools/rhs_once.py --bodies 262144
step ncu_full_sym64_general 600 ncu --set full --clock-control none --import-source on -k regex:force_sym_kernel -c 1 \
	-o "$OUT/k1s_fp64_general_262144" -f $PY tools/rhs_once.py --bodies 262144 --unimass 0
step ncu_full_sym32 600 ncu --set full --clock-control none --import-source on -k regex:force_sym_kernel -c 1 \
	-o "$OUT/k1s_fp32_262144" -f $PY tools/rhs_once.py --bodies 262144 --precision 32
step ncu_full_k1_20k 600 ncu --set full --clock-control none --import-source on -k regex:force_kernel -c 1 \
	-o "$OUT/k1_fp64_general_20000" -f $PY tools/rhs_once.py --bodies 20000 --unimass 0
cat "$OUT/summary.txt"
