#!/usr/bin/env bash
# ncu --set full of the spectral / RGB eval kernels of one library build, summarised on the box.
# Usage: bash profiles/r2_ncu_eval.sh <variant name | default> <arith> [kernel regex] [prof_driver kernels]
v=$1; A=${2:-fast}; RX=${3:-eval_spec}; K=${4:-eval_spec}
O=gpurun_out; mkdir -p $O
lib=profiles/variants/$v.so; [ "$v" = default ] && lib=mitsuba3-sunsky_b200/libsunsky_b200.so
export SSK_LIB=$PWD/$lib
T=${v}_${A}_${K//,/_}
python profiles/prof_driver.py --kernels $K --reps 5 --arith $A > $O/r2_plain_$T.log 2>&1 || { echo plain run failed; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:$RX --launch-skip 1 -c 1 -f -o $O/r2_prof_$T \
    python profiles/prof_driver.py --kernels $K --reps 1 --arith $A > $O/r2_ncu_$T.log 2>&1
python profiles/ncu_summary.py $O/r2_prof_$T.ncu-rep > $O/r2_ncu_full_$T.txt 2>&1
ncu -i $O/r2_prof_$T.ncu-rep --page source --csv > $O/r2_src_$T.csv 2>/dev/null
python profiles/src_hot.py $O/r2_src_$T.csv 40 > $O/r2_hot_$T.txt 2>&1
python profiles/sass_mix.py $O/r2_src_$T.csv 16777216 > $O/r2_sassmix_$T.txt 2>&1
ncu -i $O/r2_prof_$T.ncu-rep --page details 2>/dev/null | grep -E "Registers Per|Theoretical Occ|Achieved Occ|Eligible Warps|Issued Warp|No Eligible|One or More|Active Warps|L1/TEX Hit|Bank conflicts|Executed Ipc|Local" > $O/r2_details_$T.txt
rm -f $O/r2_prof_$T.ncu-rep $O/r2_src_$T.csv
cat $O/r2_plain_$T.log $O/r2_ncu_full_$T.txt $O/r2_details_$T.txt $O/r2_sassmix_$T.txt; head -50 $O/r2_hot_$T.txt
