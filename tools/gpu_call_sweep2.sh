#!/usr/bin/env bash
# Second K1s variant sweep: occupancy / CTA-shape / unroll builds (FP64 table entries 9-16)
# against the round's default (entry 8), at 1M (equal and general masses) and at 256k.
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/sweep2
mkdir -p "$OUT"
: > "$OUT/sweep2.jsonl"
run() {	 # n unimass variant
	timeout 300 python bench.py --bodies "$1" --unimass "$2" --pair-symmetry 1 --sym-variant "$3" \
		--steps 2 --warmup 3 --no-cpu-baseline --no-energy --no-secondary 2> "$OUT/err_$1_$2_$3.log" |
		grep '^{' >> "$OUT/sweep2.jsonl"
	echo "n=$1 um=$2 v=$3 exit ${PIPESTATUS[0]}"
}
for v in 8 9 10 11 12 13 14 15 16; do run 1048576 -1 $v; done
for v in 7 8 5 10 12 15; do run 1048576 0 $v; done
for v in 8 10 12 13 15 16; do run 262144 -1 $v; done
