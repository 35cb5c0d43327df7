#!/bin/bash
# Pre-pass block size (envs per block) and a 120-register step kernel (leaves room for the pre-pass blocks beside
# eight resident step blocks), all with two us per trip of the cold loop.  Variant libraries: sparc_b200/_lib/var/.
run() { python bench.py --no-cpu-baseline --steps 60 $1 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$2 | $1 |', '%.4f ms' % d['ms_per_step'], '%.3e' % d['value'])"; }
for rep in 1 2; do
for cfg in "--envs-per-gpu 4096" "--envs-per-gpu 16384" "--config cfg3"; do
  [ "$rep" = 2 ] && [ "$cfg" = "--config cfg3" ] && continue
  for v in u2 u2mt128 u2mt256 u2mt512 u2r120 u2r120mt128; do export SPARC_LIB_PATH=$PWD/sparc_b200/_lib/var/$v.so; run "$cfg" $v; done
done; done
