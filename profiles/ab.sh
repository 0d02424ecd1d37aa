#!/usr/bin/env bash
# A/B of tuning builds: bash profiles/ab.sh "<kernels>" lib1.so lib2.so ...   (default lib first)
K=$1; shift
for L in "" "$@"; do echo "#### ${L:-default}"; for A in fast exact; do echo "== $A"; SSK_LIB=${L:+$PWD/$L} python profiles/prof_driver.py --kernels $K --reps 30 --arith $A; done; done 2>&1 | tee gpurun_out/ab.log
