#!/usr/bin/env bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
K=sample_rgb,sample_sky,sample_spec
for A in fast exact; do echo "== sorted $A"; python profiles/prof_driver.py --kernels $K --reps 30 --arith $A; echo "== unsorted $A"; SSK_SAMPLE_NOSORT=1 python profiles/prof_driver.py --kernels $K --reps 30 --arith $A; done 2>&1 | tee gpurun_out/sort_ab.log
bash profiles/e2e_tune.sh
