#!/usr/bin/env bash
# Multi-GPU call (gpurun --gpus G): parity under torchrun, bench lines with per-rank traces,
# the C++ driver's weak-scaling / async-output check.   usage: gpu_call_multi.sh G [quick]
set -u
cd "$(dirname "$0")/.."
G=${1:-2}
OUT=gpurun_out/multi$G
mkdir -p "$OUT"
PY=python
step() { local name=$1 secs=$2; shift 2; echo "== $name" | tee -a "$OUT/summary.txt"
	timeout "$secs" "$@" > "$OUT/$name.log" 2>&1; local rc=$?; echo "   exit $rc" | tee -a "$OUT/summary.txt"; return $rc; }
TR="$PY -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1"
bench_line() { local name=$1 port=$2 secs=$3; shift 3
	step "$name" "$secs" $TR --master-port $port bench.py --gpus $G "$@" && grep '^{' "$OUT/$name.log" | tail -1 > "$OUT/$name.json"; }
nvidia-smi --query-gpu=index,name,clocks.max.sm --format=csv > "$OUT/gpus.txt" 2>&1
if [ "$G" = 2 ]; then
	step pytest_multi 1200 $PY -m pytest tests/test_gpu_multi.py tests/test_zzz_gpu_pair_symmetry.py::test_sym_two_ranks_match_oracle -m gpu -q
fi
for sym in auto 0; do
	NBODY_PAIR_SYM=$sym step multi_check_sym_$sym 600 $TR --master-port 29541 tests/tools/multi_check.py 6000
done
[ "$G" -le 4 ] && NBODY_P2P=1 step multi_check_p2p 600 $TR --master-port 29542 tests/tools/multi_check.py 6000
bench_line bench_1m 29551 600 --steps 5 --warmup 3 --trace-out "$OUT/trace_1m.json"
bench_line bench_256k 29552 400 --bodies 262144 --steps 5 --warmup 3 --trace-out "$OUT/trace_256k.json"
bench_line bench_1m_fp32 29553 400 --precision 32 --steps 5 --warmup 3
[ "${2:-}" = full ] && bench_line bench_1m_k1 29554 600 --pair-symmetry 0 --steps 2 --warmup 3 --no-secondary --trace-out "$OUT/trace_1m_k1.json"
step driver_weak 1500 $PY tools/driver_check.py 524288 $G 32
cp /dev/shm/nbody_driver_check/*.json "$OUT/" 2>/dev/null
cat "$OUT/summary.txt"
