#!/usr/bin/env bash
# One compute-sanitizer tool per GPU call (the profiling recipe allows one tool per gpurun call) over tools/sanitize_driver.py.
#   tools/sanitize.sh memcheck|initcheck|racecheck|synccheck [parts...]      logs -> gpurun_out/sanitizer_<tool>_<part>.log
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
tool=${1:-memcheck}; shift || true
parts=${*:-small fp64 emu surface}
for part in $parts; do
  log=gpurun_out/sanitizer_${tool}_${part}.log
  timeout ${SANITIZE_TIMEOUT:-600} $CS --tool $tool --print-limit 20 python tools/sanitize_driver.py $part > $log 2>&1
  echo "$tool $part: exit $? | $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $log | tail -1)"
done
