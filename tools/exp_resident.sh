#!/bin/bash
# Resident blocks per SM of the W = 1 step kernel capped with unused dynamic shared memory (library built with
# -DSPARC_EXPERIMENT_KNOBS -DSPARC_QUICK into sparc_b200/_lib/var/knobs.so): 8 (none) / 7 / 6 / 5 blocks of 2 warps.
export SPARC_LIB_PATH=$PWD/sparc_b200/_lib/var/knobs.so
for dyn in 0 29000 35000 42000; do
  for envs in 4096 16384; do
    SPARC_EXP_DYNSMEM=$dyn python bench.py --no-cpu-baseline --steps 60 --envs-per-gpu $envs 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('dynsmem $dyn envs $envs', '%.4f ms' % d['ms_per_step'], '%.3e' % d['value'])"
  done
done
