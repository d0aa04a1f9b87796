#!/bin/bash
# usage: ab.sh label [ENV=VAL ...]   -> one line per config
label=$1; shift
env "$@" python bench.py --steps 30 --warmup 30 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('$label', round(d['value']/1e6,2), 'M  kernel_ms', round(d['roofline']['kernel_ms_avg'],3))
except Exception as e: print('$label', 'FAILED', e)"
