#!/bin/bash
# Runs every group of scripts/sanitize_cases.py under compute-sanitizer memcheck / synccheck / racecheck; logs go to
# gpurun_out/sanitizer/ (copy the summaries to profiles/).  Each run is bounded by its own timeout.
mkdir -p gpurun_out/sanitizer
for tool in memcheck synccheck racecheck; do
  for g in decoder_simt decoder_tc encoder; do
    log=gpurun_out/sanitizer/${tool}_${g}.log
    t0=$SECONDS
    timeout ${SAN_TIMEOUT:-420} compute-sanitizer --tool $tool --log-file $log \
      python scripts/sanitize_cases.py $g > gpurun_out/sanitizer/${tool}_${g}.out 2>&1
    echo "$tool $g rc=$? $((SECONDS - t0)) s :: $(grep 'ERROR SUMMARY' $log 2>/dev/null | tail -1)"
  done
done
