#!/usr/bin/env bash
# Last call of round 2: the whole GPU suite, smoke(), the plain bench line on the final tree, and a
# full ncu capture of the fixed-point build at the headline size.
#   gpurun --timeout 260 -- 'bash tools/gpu_call_final4.sh'
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/final4
mkdir -p "$OUT"
PY=python
DEADLINE=${DEADLINE:-235}
step() { local name=$1 secs=$2; shift 2; echo "== $name" | tee -a "$OUT/summary.txt"
	local t0=$SECONDS left=$((DEADLINE - SECONDS))
	[ "$left" -lt 15 ] && { echo "   skipped: out of time" | tee -a "$OUT/summary.txt"; return 1; }
	[ "$secs" -gt "$left" ] && secs=$left
	timeout "$secs" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "   exit $rc  ($((SECONDS - t0)) s)" | tee -a "$OUT/summary.txt"; return $rc; }
step pytest_gpu 200 $PY -m pytest tests -m gpu -q -x
step smoke 60 $PY -c "import __graft_entry__ as g; g.smoke()"
step bench_plain 200 $PY bench.py && grep '^{' "$OUT/bench_plain.log" | tail -1 > "$OUT/bench_plain.json"
step ncu_det_1m 120 ncu --set full --clock-control none --import-source on -k regex:force_sym_kernel -s 1 -c 1 \
	-o "$OUT/k1s_fx_fp64_equal_1M" -f $PY bench.py --deterministic 1 --steps 1 --warmup 1 \
	--no-cpu-baseline --no-secondary --no-energy --no-parity
[ -f "$OUT/k1s_fx_fp64_equal_1M.ncu-rep" ] && $PY tools/ncu_summary.py "$OUT/k1s_fx_fp64_equal_1M.ncu-rep" \
	"$OUT/ncu_k1s_fx_fp64_equal_1M.csv" > /dev/null 2>&1
rm -f "$OUT"/*.ncu-rep
cat "$OUT/summary.txt"
