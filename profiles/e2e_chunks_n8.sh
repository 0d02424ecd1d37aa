#!/usr/bin/env bash
# e2e through ssk_eval_host at 8 GPUs for two chunk sizes of the staging pipeline (bench.py --no-extra)
for c in 19 21; do SSK_HOST_CHUNK_LOG2=$c python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2953$((c-18)) bench.py --gpus 8 --steps 20 --warmup 5 --no-extra 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('chunk $c', round(d['value'],1), round(e['value'],4), round(e['ms_per_step'],3), round(e['copy_ceiling_ms'],3), round(e['frac'],4))"; done
