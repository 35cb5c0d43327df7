#!/bin/bash
# usage: tools/exp_occ.sh — A/B of library variants (sparc_b200/_lib/var/*.so + the default) at several env counts
for lib in default sparc_b200/_lib/var/*.so; do
  for E in ${EXP_ENVS:-4096 4144 8192 16384}; do
    if [ "$lib" = default ]; then unset SPARC_LIB_PATH; else export SPARC_LIB_PATH=$PWD/$lib; fi
    python bench.py --steps ${EXP_STEPS:-10} --warmup 3 --no-cpu-baseline --envs-per-gpu $E 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$lib', $E, '%.3e'%d['value'], 'ms/step %.4f'%d['ms_per_step'], 'e2e %.3e'%d['e2e']['value'], 'sm_mhz', d['clocks']['sm_mhz'])"
  done
done
