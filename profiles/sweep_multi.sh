#!/bin/bash
# usage: profiles/sweep_multi.sh NGPU out.jsonl size1 size2 ...   : one weak-scaling bench line per per-GPU env count (config 5)
n=$1; out=$2; shift 2
for e in "$@"; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $n \
    --envs $e --steps 100 --warmup 25 --no-cpu --no-e2e --strong-total 0 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print(json.dumps({'n_gpus': d['n_gpus'], 'envs_per_gpu': d['config']['envs_per_gpu'], 'envs_total': d['config']['envs_total'], 'value': d['value'], 'ms_per_step': d['ms_per_step']}))" >> $out
done
