#!/bin/bash
# compute-sanitizer over the evaluation's launches: memcheck, then racecheck (shared-memory hazards
# of the specialised kernels' parking slots and the gathers' block reductions).  Summaries go to profiles/.
O=gpurun_out/${1:-sanitize}; mkdir -p $O
for tool in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 --log-file $O/$tool.log python tools/sanitize_eval.py 300 > $O/$tool.out 2>&1
  echo "$tool rc=$?"; tail -n 3 $O/$tool.out; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|error" $O/$tool.log | tail -n 5
done
