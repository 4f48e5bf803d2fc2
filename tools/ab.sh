#!/bin/bash
# A/B of kernel variants built by tools/build_variant.sh: prints the per-locus kernel time of each on the bench workload.
cd "$(dirname "$0")/.."
for v in "$@"; do
  echo "== $v"
  GUAC_B200_LIB=$PWD/build_variants/libguac_$v.so python tools/trace_run.py 2>&1 | grep "^step" | tail -2
done
