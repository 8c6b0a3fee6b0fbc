#!/usr/bin/env bash
# compute-sanitizer passes over tools/sanitize_driver.py (run on the GPU box).  Logs -> gpurun_out/sanitizer_*.log
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
for pair in memcheck:small memcheck:fp64 memcheck:emu memcheck:surface initcheck:small initcheck:emu racecheck:small racecheck:fp64 racecheck:emu synccheck:emu; do
  tool=${pair%%:*}; part=${pair##*:}
  log=gpurun_out/sanitizer_${tool}_${part}.log
  timeout ${SANITIZE_TIMEOUT:-420} $CS --tool $tool --print-limit 20 python tools/sanitize_driver.py $part > $log 2>&1
  echo "$tool $part: exit $? | $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $log | tail -1)"
done
