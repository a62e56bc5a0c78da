#!/usr/bin/env bash
# compute-sanitizer over every kernel of the library at small N (SURVEY.md section 5 plan):
# memcheck, racecheck (shared-memory hazards: the bulk-copy ring, the ticket hand-off, K1s's
# strip-commit sums), synccheck.  Logs land in gpurun_out/sanitize/; copy the summaries to profiles/.
#   gpurun --timeout 900 -- 'bash tools/sanitize.sh'
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/sanitize
mkdir -p "$OUT"
CS=${CS:-/usr/local/cuda/bin/compute-sanitizer}
for tool in memcheck racecheck synccheck; do
	extra=""
	[ "$tool" = racecheck ] && extra="--racecheck-report all"
	echo "== $tool" | tee -a "$OUT/summary.txt"
	timeout ${SAN_TIMEOUT:-600} $CS --tool $tool $extra --print-limit 40 --log-file "$OUT/$tool.log" \
		python tools/sanitize_target.py ${SAN_SIZES:-257 3000} > "$OUT/$tool.stdout" 2>&1
	echo "   exit $? ; $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' "$OUT/$tool.log" | tail -1)" | tee -a "$OUT/summary.txt"
	tail -2 "$OUT/$tool.stdout" >> "$OUT/summary.txt"
done
