#!/bin/bash
# usage: profiles/sweep_env.sh VAR v1 v2 ... -- <bench args>   : one bench.py line per value of the environment variable VAR
var=$1; shift
vals=()
while [ "$1" != "--" ] && [ $# -gt 0 ]; do vals+=("$1"); shift; done
shift
for v in "${vals[@]}"; do
  env $var=$v timeout 200 python bench.py --no-cpu --no-e2e "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$var=$v', round(d['value']/1e6,2), 'Msteps/s', round(d['ms_per_step'],4), 'ms', {k: round(x*1e3,1) for k,x in d['kernel_ms_per_launch'].items()})"
done
