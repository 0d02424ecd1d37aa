#!/usr/bin/env bash
# DEBUG AID (container only): builds libhostdbg.so, the host execution of the per-lane model.
cd "$(dirname "$0")" && /usr/local/cuda/bin/nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -fmad=false -Xcompiler -fPIC,-ffp-contract=off,-mfma \
  -DSSK_COUNT_FALLBACK --shared hostdbg.cu ../../mitsuba3-sunsky_b200/csrc/ssk_host.cu -I../../include -o libhostdbg.so
