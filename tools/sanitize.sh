#!/bin/bash
# compute-sanitizer over one small case per kernel (tools/sanitize_cases.py); SURVEY.md 5 "race detection / sanitizers".
# usage: tools/sanitize.sh memcheck|racecheck|synccheck [logfile]     -- ONE tool per call (and per gpurun call: running
# several tools back to back on one GPU has left the device unusable on this pool, see B200_PROFILING.md)
set -u
tool=${1:-memcheck}
log=${2:-gpurun_out/sanitize_$tool.log}
cd "$(dirname "$0")/.."
python tools/sanitize_cases.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool "$tool" --print-limit 20 python tools/sanitize_cases.py > "$log" 2>&1
rc=$?
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize_cases" "$log" | tail -5
exit $rc
