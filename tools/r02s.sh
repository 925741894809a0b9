#!/bin/bash
for v in lib_old lib lib_s3; do M3R_LIB_DIR=$v timeout 300 python tools/ab_solve.py 40,160 48,200 64,300 128,500 256,1000 2>&1 | grep "GN x10"; done
