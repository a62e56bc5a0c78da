#!/usr/bin/env bash
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/call3
mkdir -p "$OUT"
timeout 900 python tools/sym_crossover.py --sizes 2048,4096,8192,12288,16384,20000,24576,32768,40960,49152,65536,98304,131072 \
	--variants64=-1,8,7,5,3 --variants32=-1,1,3,2 > "$OUT/crossover.log" 2>&1
echo "crossover exit $?"
grep '^{' "$OUT/crossover.log" > "$OUT/crossover.jsonl"
