#!/usr/bin/env bash
# A/B of kernel tuning builds (make -C mitsuba3-sunsky_b200/csrc variant ...): speed through prof_driver (CUDA events),
# accuracy through the FAST-mode tolerance tests. Usage on the GPU box: bash profiles/r2_variants.sh name1 name2 ...
O=gpurun_out; mkdir -p $O
for v in "$@"; do
  lib=profiles/variants/$v.so; [ "$v" = default ] && lib=mitsuba3-sunsky_b200/libsunsky_b200.so
  echo "== $v" | tee -a $O/r2_variants.log
  SSK_LIB=$PWD/$lib python profiles/prof_driver.py --kernels ${KERNELS:-eval_spec,eval_rgb,envmap_spec,sample_spec,wavelengths} --reps 20 --arith fast 2>&1 | tee -a $O/r2_variants.log
  SSK_LIB=$PWD/$lib python -m pytest tests/test_parity_gpu.py -q -x -s -k "fast_mode" 2>&1 | grep -E "FAST vs EXACT|passed|failed|Error" | tee -a $O/r2_variants.log
done
