#!/bin/bash
# Parity (tools/check_c2.py) and kernel time (tools/time_trunc.py) of every library variant present
# (gpsig_b200/libv_*.so, built with GPSIG_LIB_NAME / GPSIG_EXTRA_DEFS), next to the default build.
n=${1:-1000}
echo "== default"; timeout 300 python tools/check_c2.py 2>&1 | grep -c " ok"; timeout 200 python tools/time_trunc.py $n 2>&1 | tail -2
for f in gpsig_b200/libv_*.so; do
  b=$(basename $f)
  echo "== $b"
  GPSIG_LIB_NAME=$b timeout 300 python tools/check_c2.py 2>&1 | grep "FAIL\|Error\|error" | head -5
  GPSIG_LIB_NAME=$b timeout 200 python tools/time_trunc.py $n 2>&1 | tail -2
done
