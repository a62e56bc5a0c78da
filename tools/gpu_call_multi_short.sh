#!/usr/bin/env bash
# Short multi-GPU call: usage gpu_call_multi_short.sh G "steps to run" (names: check b1m b256k bfp32 weak)
set -u
cd "$(dirname "$0")/.."
G=${1:-4}; WHAT=${2:-"check b1m b256k weak"}
OUT=gpurun_out/multi${G}b
mkdir -p "$OUT"
PY=python
step() { local name=$1 secs=$2; shift 2; echo "== $name" | tee -a "$OUT/summary.txt"
	timeout "$secs" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "   exit $rc" | tee -a "$OUT/summary.txt"; return $rc; }
TR="$PY -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1"
bench_line() { local name=$1 port=$2 secs=$3; shift 3
	step "$name" "$secs" $TR --master-port $port bench.py --gpus $G "$@" && grep '^{' "$OUT/$name.log" | tail -1 > "$OUT/$name.json"; }
for w in $WHAT; do case $w in
	check) NBODY_PAIR_SYM=auto step multi_check_sym_auto 600 $TR --master-port 29541 tests/tools/multi_check.py 6000;;
	b1m) bench_line bench_1m 29551 600 --steps 5 --warmup 3 --trace-out "$OUT/trace_1m.json";;
	b256k) bench_line bench_256k 29552 400 --bodies 262144 --steps 10 --warmup 3 --trace-out "$OUT/trace_256k.json";;
	bfp32) bench_line bench_1m_fp32 29553 400 --precision 32 --steps 5 --warmup 3;;
	weak) step driver_weak 1500 $PY tools/driver_check.py 524288 $G 32;;
esac; done
cat "$OUT/summary.txt"
