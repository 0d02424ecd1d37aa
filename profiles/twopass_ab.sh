#!/usr/bin/env bash
# spectral sample_direction: fused kernel vs sampler pass + weight pass (sorted / unsorted first pass)
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
K=sample_spec,sample_spec_sky
for A in fast exact; do
  echo "== two-pass sorted $A"; python profiles/prof_driver.py --kernels $K --reps 30 --arith $A
  echo "== two-pass unsorted $A"; SSK_SAMPLE_2P_NOSORT=1 python profiles/prof_driver.py --kernels $K --reps 30 --arith $A
  echo "== fused $A"; SSK_SAMPLE_FUSED=1 python profiles/prof_driver.py --kernels $K --reps 30 --arith $A
done 2>&1 | tee gpurun_out/r1_twopass_ab.log
# spectral sample_ray: fused vs geometry pass + wavelength/weight pass
for A in fast exact; do
  echo "== ray two-pass $A"; python profiles/prof_driver.py --kernels ray_spec --reps 30 --arith $A
  echo "== ray fused $A"; SSK_SAMPLE_FUSED=1 python profiles/prof_driver.py --kernels ray_spec --reps 30 --arith $A
done 2>&1 | tee -a gpurun_out/r1_twopass_ab.log
