#!/bin/bash
# compute-sanitizer record: memcheck, racecheck, synccheck, initcheck over scripts/sanitize_run.py
# (update in both modes + force_search, the ikd-Tree golden map sequence, a 2-rank peer exchange on one GPU).
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
for tool in memcheck racecheck synccheck initcheck; do
  for what in update map peer; do
    echo "=== $tool / $what" | tee -a gpurun_out/sanitizer.txt
    timeout 1500 $CS --tool $tool --print-limit 20 python scripts/sanitize_run.py $what > gpurun_out/san_${tool}_${what}.log 2>&1
    echo "rc=$?" | tee -a gpurun_out/sanitizer.txt
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|Uninitialized|ok$|Error|error" gpurun_out/san_${tool}_${what}.log | sort | uniq -c | sort -rn | head -12 | tee -a gpurun_out/sanitizer.txt
  done
done
