#!/usr/bin/env bash
# ncu --set full of selected kernels: bash profiles/ncu_one.sh <kernels csv> <arith> <tag>
K=${1:-eval_spec}; A=${2:-fast}; T=${3:-one}; O=gpurun_out; mkdir -p $O
python profiles/prof_driver.py --kernels $K --reps 1 --arith $A > $O/plain_$T.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'eval_|pdf_|sample_|envmap_kernel' -c 12 -f -o $O/prof_$T \
    python profiles/prof_driver.py --kernels $K --reps 1 --arith $A > $O/ncu_$T.log 2>&1
echo "rc=$?"; tail -3 $O/ncu_$T.log
