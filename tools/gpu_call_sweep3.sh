#!/usr/bin/env bash
# 64-thread K1s builds (FP64 entries 9, 10; FP32 entry 5) against the incumbents, per size.
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/sweep3
mkdir -p "$OUT"
timeout 800 python tools/sym_crossover.py \
	--sizes 4096,8192,16384,20000,32768,65536,131072,262144,524288,1048576 \
	--variants64=-1,3,5,7,8,9,10 --variants32=-1,1,3,4,5 > "$OUT/crossover.log" 2>&1
echo "crossover exit $?"
grep '^{' "$OUT/crossover.log" > "$OUT/crossover.jsonl"
