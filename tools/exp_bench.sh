#!/bin/bash
# usage: tools/exp_bench.sh label [ENV=VAL ...] -- runs bench at several env counts, prints value
label=$1; shift
for E in ${EXP_ENVS:-4096 16384}; do
  env "$@" python bench.py --steps ${EXP_STEPS:-10} --warmup 3 --no-cpu-baseline --envs-per-gpu $E 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$label', $E, '%.3e'%d['value'], 'ms/step %.4f'%d['ms_per_step'], 'e2e %.3e'%d['e2e']['value'], 'sm_mhz', d['clocks']['sm_mhz'])"
done
