#!/usr/bin/env bash
# Option "deterministic" on several GPUs (integer reduce-scatter of the fixed-point limbs over NCCL):
# every rank against the oracle + bit-identical repeats, then the mode's bench lines.
#   gpurun --gpus 2 --timeout 190 -- 'bash tools/gpu_call_det_multi.sh 2'
set -u
cd "$(dirname "$0")/.."
G=${1:-2}
OUT=gpurun_out/det_multi${G}
mkdir -p "$OUT"
PY=python
DEADLINE=${DEADLINE:-170}
step() { local name=$1 secs=$2; shift 2; echo "== $name" | tee -a "$OUT/summary.txt"
	local t0=$SECONDS left=$((DEADLINE - SECONDS))
	[ "$left" -lt 15 ] && { echo "   skipped: out of time" | tee -a "$OUT/summary.txt"; return 1; }
	[ "$secs" -gt "$left" ] && secs=$left
	timeout "$secs" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "   exit $rc  ($((SECONDS - t0)) s)" | tee -a "$OUT/summary.txt"; return $rc; }
TR="$PY -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1"
bench_line() { local name=$1 port=$2 secs=$3; shift 3
	step "$name" "$secs" $TR --master-port $port bench.py --gpus $G "$@" && grep '^{' "$OUT/$name.log" | tail -1 > "$OUT/$name.json"; }
NBODY_DETERMINISTIC=1 NBODY_PAIR_SYM=auto step multi_check_det 120 $TR --master-port 29561 tests/tools/multi_check.py 6000
bench_line bench_det_256k 29562 60 --deterministic 1 --bodies 262144 --steps 5 --warmup 3 --no-energy --no-secondary
bench_line bench_det_1m 29563 90 --deterministic 1 --steps 2 --warmup 2 --no-energy --no-secondary
step pytest_multi_det 90 $PY -m pytest tests/test_gpu_multi.py -m gpu -q -x -k deterministic
cat "$OUT/summary.txt"
