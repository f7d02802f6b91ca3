#!/bin/bash
# First-contact GPU diagnostics: each pytest file in its own process with a timeout so a
# hung kernel cannot take the whole call down.  Logs land in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/smi.txt 2>&1
for t in "$@"; do
  name=$(basename "$t" .py)
  timeout 300 python -m pytest "$t" -m gpu -q --tb=short --no-header -p no:cacheprovider > "gpurun_out/${name}.log" 2>&1
  echo "$name exit $?" | tee -a gpurun_out/summary.txt
  tail -5 "gpurun_out/${name}.log"
done
