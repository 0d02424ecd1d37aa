#!/usr/bin/env bash
# One gpurun call of round 1: GPU parity tests, smoke, both bench arms, then ncu (launch list + full capture).
# Usage (from the repo root, on the GPU box): bash profiles/run_gpu_r1.sh
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.csv 2>&1
lscpu | head -20 > $O/lscpu.txt 2>&1
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "ref rc=$?"
python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
python bench.py --arith exact --no-extra --steps 100 > $O/bench_exact.json 2> $O/bench_exact.err; echo "bench exact rc=$?"
# launch list of the bench command itself (its kernel share of the step is what must agree with the live timing)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench.csv \
    python bench.py --steps 20 --warmup 3 --no-extra > $O/ncu_launches_bench.log 2>&1; echo "bench launch list rc=$?"
K="setup,eval_spec,eval_rgb,pdf,sample_rgb,sample_sky,sample_spec,envmap,wavelengths,ray_rgb,ray_spec"
for A in fast exact; do
python profiles/prof_driver.py --kernels $K --reps 3 --arith $A > $O/prof_plain_$A.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_$A.csv \
    python profiles/prof_driver.py --kernels $K --reps 3 --arith $A > $O/ncu_launches_$A.log 2>&1
echo "launch list $A rc=$?"
done
# full captures: summarised on the box (gpurun_out/ is capped at 64 MiB, a 40-kernel .ncu-rep is larger)
KF="eval_spec,eval_rgb,pdf,sample_sky"
python profiles/prof_driver.py --kernels $KF --reps 1 --arith fast > $O/prof_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'eval_|pdf_|sample_' -c 8 -f -o $O/prof_fast \
    python profiles/prof_driver.py --kernels $KF --reps 1 --arith fast > $O/ncu_full_fast.log 2>&1
echo "full fast rc=$?"
python profiles/ncu_summary.py $O/prof_fast.ncu-rep > $O/ncu_full_fast.txt 2>&1
ncu -i $O/prof_fast.ncu-rep --page source --csv --kernel-name regex:eval_spec > $O/src_eval_spec_fast.csv 2>/dev/null
python profiles/src_hot.py $O/src_eval_spec_fast.csv 30 > $O/hot_eval_spec_fast.txt 2>&1
KS="setup,sample_rgb,sample_spec,wavelengths,ray_rgb,ray_spec,envmap"
ncu --set full --clock-control none -k regex:'sample_|setup_|envmap_' --launch-skip 10 -c 24 -f -o $O/prof_fast2 \
    python profiles/prof_driver.py --kernels $KS --reps 1 --arith fast > $O/ncu_full_fast2.log 2>&1
echo "full fast2 rc=$?"
python profiles/ncu_summary.py $O/prof_fast2.ncu-rep > $O/ncu_full_fast2.txt 2>&1; rm -f $O/prof_fast2.ncu-rep
ncu --set full --clock-control none -k regex:'eval_|pdf_|sample_' -c 12 -f -o $O/prof_exact \
    python profiles/prof_driver.py --kernels eval_spec,eval_rgb,pdf,sample_rgb,sample_sky,ray_spec --reps 1 --arith exact > $O/ncu_full_exact.log 2>&1
echo "full exact rc=$?"
python profiles/ncu_summary.py $O/prof_exact.ncu-rep > $O/ncu_full_exact.txt 2>&1; rm -f $O/prof_exact.ncu-rep
du -sh $O; ls -la $O
