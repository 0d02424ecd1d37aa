#!/usr/bin/env bash
# Quick GPU iteration: parity tests, smoke, then CUDA-event timings of every kernel in both arithmetic modes.
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee $O/quick_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5 | tee $O/quick_smoke.log
K="setup,eval_spec,eval_rgb,pdf,sample_rgb,sample_sky,sample_spec,envmap,wavelengths,ray_rgb,ray_spec"
for A in fast exact; do echo "== $A"; python profiles/prof_driver.py --kernels $K --reps 20 --arith $A; done 2>&1 | tee $O/quick_prof.log
