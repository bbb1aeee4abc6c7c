#!/bin/bash
# Diagnostic: run the conv parity tests and the per-op profile with each library build under variants/*.so (git-ignored),
# then restore the default build.  Usage (GPU box): bash scripts/variants.sh [workload ...]
cd "$(dirname "$0")/.."
LIB=calodiffusion_b200/libcalodiffusion_b200.so
cp $LIB /tmp/default.so
WL="${@:-ds2}"
for v in variants/*.so; do
  echo "=== $v"
  cp $v $LIB
  timeout 600 python -m pytest tests/test_gpu_parity.py -q -x -m gpu -k "tcgen05_conv or denoise_matches_reference_golden" 2>&1 | tail -3
  for w in $WL; do timeout 300 python scripts/eval_profile.py $w 10 2>&1 | tail -1; done
done
cp /tmp/default.so $LIB
