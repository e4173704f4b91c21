#!/bin/bash
# Sweep an environment knob over refine-loop configs.  Usage: bash scripts/gpu_sweep.sh <VAR> "<values>" "<configs>"
VAR=$1; VALS=$2; CFGS=${3:-"3 5"}
for v in $VALS; do for c in $CFGS; do
export $VAR=$v
python bench.py --config $c --steps 40 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$VAR=$v cfg$c', round(d['value']), 'ms/step', round(d['ms_per_step'],3))"
done; done
