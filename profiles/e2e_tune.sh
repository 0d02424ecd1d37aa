#!/usr/bin/env bash
for c in 18 19 20 21 22; do for s in 2 3 4 6; do SSK_HOST_CHUNK_LOG2=$c SSK_HOST_STREAMS=$s python profiles/e2e_tune.py 2>&1 | tail -1; done; done | tee gpurun_out/e2e_tune.log
