#!/usr/bin/env bash
# Round-3 single-GPU call: GPU suite, default bench line, the K2 (stage_kernel) ncu capture the
# round-2 review asked for, and nbody_rhs_noalloc timed against nbody_rhs.
#   gpurun --timeout 1200 -- 'bash tools/gpu_call_r03.sh'
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/r03b
mkdir -p "$OUT"
PY=python
step() { local name=$1 secs=$2; shift 2; echo "== $name" | tee -a "$OUT/summary.txt"
	timeout "$secs" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "   exit $rc" | tee -a "$OUT/summary.txt"; return $rc; }
step pytest_gpu 900 $PY -m pytest tests -m gpu -x -q
step bench_default 400 $PY bench.py && grep '^{' "$OUT/bench_default.log" | tail -1 > "$OUT/bench_default.json"
step noalloc_timing 200 $PY tools/noalloc_timing.py
step bounds_check_noalloc 300 env NBODY_B200_LIB=parallel_comp_labs_b200/libnbody_b200_check.so $PY -m pytest tests/test_zzzzz_no_alloc.py -m gpu -x -q
if step step_once 200 $PY tools/step_once.py; then
	step ncu_k2_full 600 ncu --set full --clock-control none --import-source on -k regex:stage_kernel -c 2 \
		-o "$OUT/stage_kernel_1M" -f $PY tools/step_once.py
	step ncu_k2_summary 120 $PY tools/ncu_summary.py "$OUT/stage_kernel_1M.ncu-rep" "$OUT/r03_ncu_k2_fp64_1M.csv"
	step ncu_k2_raw 120 bash -c "ncu -i $OUT/stage_kernel_1M.ncu-rep --page raw --csv > $OUT/stage_kernel_1M_raw.csv"
	rm -f "$OUT/stage_kernel_1M.ncu-rep"
fi
cat "$OUT/summary.txt"
