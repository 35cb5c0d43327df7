#!/bin/bash
# persistent vs one-warp-per-env launches of one library at several env counts
export SPARC_LIB_PATH=$PWD/$1
for E in 2368 4096 4736 7104 16384; do
  for mode in persist nopersist; do
    if [ $mode = nopersist ]; then export SPARC_NO_PERSIST=1; else unset SPARC_NO_PERSIST; fi
    python bench.py --steps 10 --warmup 3 --no-cpu-baseline --envs-per-gpu $E 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$mode', $E, '%.3e'%d['value'], 'ms/step %.4f'%d['ms_per_step'], 'blocks', d['config']['launch']['blocks'])"
  done
done
