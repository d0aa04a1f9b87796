#!/bin/bash
# usage: ab2.sh label workload envs [ENV=VAL ...]
label=$1; wl=$2; envs=$3; shift 3
env "$@" python bench.py --steps 30 --warmup 30 --no-cpu-baseline --workload $wl --envs $envs 2>&1 | tail -1 | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('$label', '$wl', $envs, round(d['value']/1e6,3), 'M  kernel_ms', round(d['roofline']['kernel_ms_avg'],3))
except Exception as e: print('$label', 'FAILED', e)"
