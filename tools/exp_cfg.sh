#!/bin/bash
# usage: tools/exp_cfg.sh label "bench args" [ENV=VAL ...] -- one bench line, value / e2e / ms per step
label=$1; args=$2; shift 2
env "$@" python bench.py --warmup 5 --no-cpu-baseline $args 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$label', '$args', '%.3e'%d['value'], 'ms/step %.4f'%d['ms_per_step'], 'e2e %.3e'%d['e2e']['value'])"
