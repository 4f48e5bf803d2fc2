#!/bin/bash
# Rebuilds the CUDA library and fails loudly on any compiler error (so a stale .so never travels to the GPU box).
cd "$(dirname "$0")/.."
python __graft_entry__.py --force > /tmp/build_out.txt 2>&1
if grep -E "error" guacamole_b200/csrc/build.log | grep -v "DeviceError\|report_error\|guac_last_error\|_error" ; then echo "BUILD FAILED"; exit 1; fi
grep -A3 "$1" guacamole_b200/csrc/build.log | grep -E "Used|spill" | head -${2:-6}
echo "build ok"
