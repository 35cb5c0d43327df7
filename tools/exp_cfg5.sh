#!/bin/bash
# cfg5 (full-field logging) throughput for the default library and every variant under sparc_b200/_lib/var
for lib in default sparc_b200/_lib/var/*.so; do
  if [ "$lib" = default ]; then unset SPARC_LIB_PATH; else [ -f "$lib" ] || continue; export SPARC_LIB_PATH=$PWD/$lib; fi
  for cfg in "0.05 1" "0.05 10" "0.02 1" "0.02 10" "0.1 1"; do
    set -- $cfg
    python bench.py --config cfg5 --segment-len $1 --log-every $2 --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; print('$lib seg $1 every $2', 'N', d['config']['n_segments'], '%.3e env-us/s'%d['value'], '%.0f GB/s'%r['achieved'], 'frac %.3f'%r['frac'])"
  done
done
