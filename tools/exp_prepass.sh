#!/bin/bash
# Late round 2: does the pre-pass kernel's register footprint cost the step kernel a resident block per SM?
#   u2: two us per trip of the cold loop (main: four); knobs: the build before the prefetch reorder, two us per trip;
#   mt64 / mt128: pre-pass blocks of 64 / 128 threads (fewer SMs host one);  r112: step kernel compiled for 9+ blocks
#   per SM (96 registers), residency capped with unused dynamic shared memory (needs the knobs build).
run() { python bench.py --no-cpu-baseline --steps 60 --envs-per-gpu $1 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$2 envs $1', '%.4f ms' % d['ms_per_step'], '%.3e' % d['value'])"; }
for rep in 1 2; do
for envs in 4096 16384; do
  unset SPARC_LIB_PATH SPARC_EXP_DYNSMEM; run $envs main
  for v in u2 knobs mt64 mt128; do export SPARC_LIB_PATH=$PWD/sparc_b200/_lib/var/$v.so; run $envs $v; done
  export SPARC_LIB_PATH=$PWD/sparc_b200/_lib/var/r112.so
  run $envs r96-10blocks
  SPARC_EXP_DYNSMEM=21000 run $envs r96-9blocks
  SPARC_EXP_DYNSMEM=24000 run $envs r96-8blocks
done; done
