#!/usr/bin/env bash
# Kernel timings only (no tests): bash profiles/quick2.sh [kernels] [reps]
K=${1:-eval_spec,eval_rgb,pdf,sample_rgb,sample_sky,sample_spec,envmap}
for A in fast exact; do echo "== $A"; python profiles/prof_driver.py --kernels $K --reps ${2:-20} --arith $A; done 2>&1 | tee gpurun_out/quick2.log
