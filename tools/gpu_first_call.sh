#!/usr/bin/env bash
# First GPU call of a round: everything that has only been run under CPU emulation so far
# (K2g general tableaux, K1s pair-symmetric kernel) plus a fresh default bench line and the ncu
# evidence for whichever force kernel wins.  One B200; every step has its own `timeout`, ncu runs
# only after the same command has exited 0 without it, and all output lands in gpurun_out/.
#
#   gpurun --timeout 1500 -- 'bash tools/gpu_first_call.sh'
#
# Reading the result here: gpurun_out/first_call/*.log, *.json (one bench line each), *.csv.
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/first_call
mkdir -p "$OUT"
PY=python

step() {  # step <name> <seconds> <command...>: run, log, record the exit code, never abort
	local name=$1 secs=$2; shift 2
	echo "== $name" | tee -a "$OUT/summary.txt"
	timeout "$secs" "$@" > "$OUT/$name.log" 2>&1
	local rc=$?
	echo "   exit $rc" | tee -a "$OUT/summary.txt"
	return $rc
}
bench_line() {  # bench_line <name> <seconds> <bench args...>: keeps the JSON line as <name>.json
	local name=$1 secs=$2; shift 2
	step "$name" "$secs" $PY bench.py "$@" && grep '^{' "$OUT/$name.log" | tail -1 > "$OUT/$name.json"
}

nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "$OUT/gpu.txt" 2>&1

# 1. parity: the whole -m gpu suite (the tableau and pair-symmetry files sort last)
step pytest_gpu 900 $PY -m pytest tests -m gpu -x -q

# 2. default path, the line the driver will take
bench_line bench_default 400 --steps 2 --warmup 3

# 3. pair-symmetric kernel: variants, general masses, FP32 (no CPU leg, no energy: kernel numbers)
for v in 0 1 2 3 4 5 6 7 8; do
	bench_line "bench_sym_v$v" 200 --pair-symmetry 1 --sym-variant $v --steps 2 --warmup 3 \
		--no-cpu-baseline --no-energy
done
for v in 1 5 6 7 8; do
	bench_line "bench_sym_general_v$v" 200 --pair-symmetry 1 --sym-variant $v --unimass 0 --steps 2 --warmup 3 \
		--no-cpu-baseline --no-energy
done
bench_line bench_default_general 200 --unimass 0 --steps 2 --warmup 3 --no-cpu-baseline --no-energy
for v in 0 1 2 3 4; do
	bench_line "bench_sym_fp32_v$v" 200 --pair-symmetry 1 --sym-variant $v --precision 32 --steps 2 --warmup 3 \
		--no-cpu-baseline --no-energy
done
bench_line bench_default_fp32 200 --precision 32 --steps 2 --warmup 3 --no-cpu-baseline --no-energy
bench_line bench_default_20k 200 --bodies 20000 --dist uniform --steps 10 --warmup 3 --no-cpu-baseline --no-energy
bench_line bench_sym_20k 200 --pair-symmetry 1 --bodies 20000 --dist uniform --steps 10 --warmup 3 --no-cpu-baseline --no-energy
step fma_peak 120 $PY tools/fma_peak.py && cp "$OUT/fma_peak.log" "$OUT/fma_peak.json"
bench_line bench_sym_256k 200 --pair-symmetry 1 --bodies 262144 --steps 3 --warmup 3 --no-cpu-baseline --no-energy
bench_line bench_default_256k 200 --bodies 262144 --steps 3 --warmup 3 --no-cpu-baseline --no-energy

# 4. ncu: launch list of the default bench command, then one full capture per force kernel at a
#    size that keeps the replays short (N = 131072: ~16 ms per launch)
if [ -s "$OUT/bench_default.json" ]; then
	step ncu_launches 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
		--log-file "$OUT/launches.csv" $PY bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-energy
	step ncu_full_default 600 ncu --set full --clock-control none --import-source on \
		-k regex:force_kernel -c 1 -o "$OUT/force_kernel_131072" -f \
		$PY tools/rhs_once.py --bodies 131072
fi
if [ -s "$OUT/bench_sym_v0.json" ]; then
	step ncu_full_sym 600 ncu --set full --clock-control none --import-source on \
		-k regex:force_sym_kernel -c 1 -o "$OUT/force_sym_kernel_131072" -f \
		$PY tools/rhs_once.py --bodies 131072 --pair-symmetry 1
fi
for cfg in "fp32 --precision 32" "general --unimass 0"; do
	set -- $cfg; nm=$1; shift
	step ncu_full_default_$nm 600 ncu --set full --clock-control none --import-source on \
		-k regex:force_kernel -c 1 -o "$OUT/force_kernel_131072_$nm" -f \
		$PY tools/rhs_once.py --bodies 131072 "$@"
	step ncu_full_sym_$nm 600 ncu --set full --clock-control none --import-source on \
		-k regex:force_sym_kernel -c 1 -o "$OUT/force_sym_kernel_131072_$nm" -f \
		$PY tools/rhs_once.py --bodies 131072 --pair-symmetry 1 "$@"
done
# 5. every kernel once at small N against the bounds-checked build (device-side asserts,
#    libnbody_b200_check.so).  compute-sanitizer is closed on this pool (profiles/r02_sanitizer.md:
#    it exits 86 and runs under it have left GPUs needing a reset), so tools/sanitize.sh is NOT
#    part of the standard call any more; NBODY_RUN_SANITIZER=1 brings it back on a pool that allows it.
NBODY_B200_LIB=parallel_comp_labs_b200/libnbody_b200_check.so \
	step bounds_check_target 300 $PY tools/sanitize_target.py
if [ "${NBODY_RUN_SANITIZER:-0}" = "1" ]; then
	SAN_TIMEOUT=420 bash tools/sanitize.sh > "$OUT/sanitize.log" 2>&1
	cat gpurun_out/sanitize/summary.txt >> "$OUT/summary.txt"
fi
cat "$OUT/summary.txt"
