#!/bin/bash
# compute-sanitizer over scripts/sanitize_workload.py: memcheck, synccheck, racecheck.  Summaries -> gpurun_out/sanitizer_<tool>.txt
# (copy into profiles/ to keep).  Each tool runs under its own timeout so a hang cannot hold the box.
set -u
mkdir -p gpurun_out
SAN=${SAN:-/usr/local/cuda/bin/compute-sanitizer}
for tool in ${TOOLS:-memcheck synccheck racecheck}; do
  out=gpurun_out/sanitizer_${tool}.txt
  echo "== $tool ==" > $out
  start=$(date +%s)
  timeout ${SAN_TIMEOUT:-900} $SAN --tool $tool --print-limit 20 --error-exitcode 3 python scripts/sanitize_workload.py >> $out 2>&1
  echo "exit code $? after $(( $(date +%s) - start )) s" >> $out
  tail -4 $out
done
