#!/bin/bash
# time every library variant present (gpsig_b200/libv_*.so, built with GPSIG_LIB_NAME / GPSIG_EXTRA_DEFS) at the c2 shape
n=${1:-1000}
python tools/time_trunc.py $n 2>&1 | tail -2
for f in gpsig_b200/libv_*.so; do
  GPSIG_LIB_NAME=$(basename $f) timeout 120 python tools/time_trunc.py $n 2>&1 | tail -2
done
