#!/bin/bash
# one-warp-per-env launches of each variant library at env counts around one wave (fixed cost per launch)
export SPARC_NO_PERSIST=1
for lib in sparc_b200/_lib/var/*.so; do
  export SPARC_LIB_PATH=$PWD/$lib
  for E in ${EXP_ENVS:-592 1184 2368 4736 9472}; do
    for pdl in on off; do
      if [ $pdl = off ]; then export SPARC_NO_PDL=1; else unset SPARC_NO_PDL; fi
      python bench.py --steps 10 --warmup 3 --no-cpu-baseline --envs-per-gpu $E 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$lib pdl=$pdl', $E, 'ms/step %.4f'%d['ms_per_step'])"
    done
  done
done
