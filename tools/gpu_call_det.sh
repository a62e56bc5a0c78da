#!/usr/bin/env bash
# Option "deterministic" (fixed-point row sums) meets the hardware: its GPU tests, the default bench
# line with the deterministic secondary figure, the mode as the main line at the benchmarked sizes,
# one ncu capture of the fixed-point build, then the whole GPU suite with what time is left.
#   gpurun --timeout 600 -- 'bash tools/gpu_call_det.sh'
set -u
cd "$(dirname "$0")/.."
OUT=${OUT:-gpurun_out/det2}
mkdir -p "$OUT"
PY=python
DEADLINE=${DEADLINE:-510}	# seconds of wall clock this script may use in all (GPU budget)
step() { local name=$1 secs=$2; shift 2; echo "== $name" | tee -a "$OUT/summary.txt"
	local t0=$SECONDS left=$((DEADLINE - SECONDS))
	[ "$left" -lt 20 ] && { echo "   skipped: out of time" | tee -a "$OUT/summary.txt"; return 1; }
	[ "$secs" -gt "$left" ] && secs=$left
	timeout "$secs" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "   exit $rc  ($((SECONDS - t0)) s)" | tee -a "$OUT/summary.txt"; return $rc; }
bench_line() { local name=$1 secs=$2; shift 2
	step "$name" "$secs" $PY bench.py "$@" && grep '^{' "$OUT/$name.log" | tail -1 > "$OUT/$name.json"; }
step pytest_gpu 400 $PY -m pytest tests -m gpu -q -x
step smoke 60 $PY -c "import __graft_entry__ as g; g.smoke()"
bench_line bench_det_1m 120 --deterministic 1 --steps 2 --warmup 3 --no-cpu-baseline --no-secondary --no-energy
bench_line bench_det_general_1m 120 --deterministic 1 --unimass 0 --steps 2 --warmup 2 --no-cpu-baseline --no-secondary --no-energy
bench_line bench_det_256k 60 --deterministic 1 --bodies 262144 --steps 5 --warmup 3 --no-cpu-baseline --no-secondary --no-energy
bench_line bench_det_20k 60 --deterministic 1 --bodies 20000 --dist uniform --steps 20 --warmup 5 --no-cpu-baseline --no-secondary --no-energy
bench_line bench_det_fp32 60 --deterministic 1 --precision 32 --steps 2 --warmup 3 --no-cpu-baseline --no-secondary --no-energy
# one full capture of the fixed-point build at N = 256k (the same command has just exited 0 above)
step ncu_det_256k 150 ncu --set full --clock-control none --import-source on -k regex:force_sym_kernel -s 2 -c 1 \
	-o "$OUT/k1s_fx_fp64_equal_262144" -f $PY bench.py --deterministic 1 --bodies 262144 --steps 1 --warmup 1 \
	--no-cpu-baseline --no-secondary --no-energy --no-parity
[ -f "$OUT/k1s_fx_fp64_equal_262144.ncu-rep" ] && $PY tools/ncu_summary.py "$OUT/k1s_fx_fp64_equal_262144.ncu-rep" \
	"$OUT/ncu_k1s_fx_fp64_equal_262144.csv" > /dev/null 2>&1
bench_line bench_default 240 --steps 2 --warmup 3 --cpu-seconds 4
cat "$OUT/summary.txt"
