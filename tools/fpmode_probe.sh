#!/bin/bash
# Parity metrics and step time of the shading fp modes (run on the GPU box).
cd "$(dirname "$0")/.."
for m in "$@"; do
  MCRT_FP_MODE=$m python -c "
import mcrt_b200
from mcrt_b200 import build as b
b.build(force=True)" > /dev/null 2>&1 || { echo "BUILD FAILED: $m"; continue; }
  echo "=== fp mode $m"
  MCRT_FP_MODE=$m python -m pytest tests -m gpu -q 2>&1 | tail -4
  MCRT_FP_MODE=$m python tools/diag_pt.py knot 2>&1 | grep -E "frac ok" | tail -4
  MCRT_FP_MODE=$m python tools/diag_bdpt.py spheres 2>&1 | grep -E "depth 5" | cut -c1-60
  MCRT_FP_MODE=$m tools/variant_bench.sh ""
done
