for c in 21 20 19 18; do for st in 3 4; do SSK_HOST_CHUNK_LOG2=$c SSK_HOST_STREAMS=$st python bench.py --steps 20 --warmup 5 --no-extra 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('chunk $c streams $st', round(e['value'],4), round(e['ms_per_step'],3), round(e['copy_ceiling_ms'],3), round(e['frac'],4))"; done; done
