#!/bin/bash
python tools/time_others.py pde 2>&1 | head -1
for f in gpsig_b200/libv_*.so; do echo $f; GPSIG_LIB_NAME=$(basename $f) timeout 120 python tools/time_others.py pde 2>&1 | head -1; done
