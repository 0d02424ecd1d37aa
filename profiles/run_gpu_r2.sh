#!/usr/bin/env bash
# Round 2, one GPU: both bench arms, the ncu launch list of the bench command, and ncu --set full of the headline kernel
# (FAST and EXACT). Usage on the GPU box: bash profiles/run_gpu_r2.sh
set -u
O=gpurun_out; mkdir -p $O
python bench.py --impl reference --steps 20 --warmup 5 > $O/r2_bench_reference.json 2> $O/r2_bench_reference.err; echo "ref rc=$?"
python bench.py --steps 20 --warmup 5 > $O/r2_bench.json 2> $O/r2_bench.err; echo "bench rc=$?"
python bench.py --arith exact --no-extra --steps 50 > $O/r2_bench_exact.json 2> $O/r2_bench_exact.err; echo "bench exact rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_launches_bench.csv \
    python bench.py --steps 20 --warmup 3 --no-extra > $O/r2_ncu_launches_bench.log 2>&1; echo "launch list rc=$?"
bash profiles/r2_ncu_eval.sh default fast eval_spec eval_spec > $O/r2_eval_spec_fast.txt 2>&1
bash profiles/r2_ncu_eval.sh default exact eval_spec eval_spec > $O/r2_eval_spec_exact.txt 2>&1
bash profiles/r2_ncu_eval.sh default fast eval_rgb eval_rgb > $O/r2_eval_rgb_fast.txt 2>&1
bash profiles/r2_ncu_eval.sh default fast "sample_" sample_rgb > $O/r2_sample_rgb_fast.txt 2>&1
ls -la $O | head -40
