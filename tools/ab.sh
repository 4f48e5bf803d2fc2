#!/bin/bash
# A/B of kernel variants built by tools/build_variant.sh: prints the per-locus kernel time of each on the bench workload.
# usage: tools/ab.sh [-i IMPL] variant...
cd "$(dirname "$0")/.."
impl=""
if [ "$1" = "-i" ]; then impl=$2; shift 2; fi
for v in "$@"; do
  echo "== $v"
  GUAC_B200_LIB=$PWD/build_variants/libguac_$v.so python tools/trace_run.py $impl 2>&1 | grep "^step" | tail -1
done
