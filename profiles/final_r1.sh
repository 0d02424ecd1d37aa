#!/usr/bin/env bash
# Round-1 closing evidence: GPU suite, smoke, JVP timing, the three bench lines, per-entry timings in both modes.
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee $O/final_tests.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee -a $O/final_tests.log
for A in fast exact; do echo "== $A"; python profiles/prof_driver.py --kernels jvp_rgb,jvp_spec --reps 20 --arith $A 2>&1 | tail -2; done | tee $O/r1_jvp_time.log
python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "ref rc=$?"
python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
python bench.py --arith exact --no-extra --steps 100 > $O/bench_exact.json 2> $O/bench_exact.err; echo "exact rc=$?"
K=eval_spec,eval_rgb,pdf,sample_rgb,sample_sky,sample_spec,sample_spec_sky,wavelengths,ray_rgb,ray_spec,envmap,envmap_spec,setup
for A in fast exact; do echo "== $A"; python profiles/prof_driver.py --kernels $K --reps 30 --arith $A 2>&1 | tail -15; done | tee $O/r1_entries.log
