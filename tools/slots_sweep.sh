#!/bin/bash
# pipeline throughput against the number of DM trials in flight per worker (PS_SLOTS)
for s in 1 2 3 4; do
  PS_SLOTS=$s python bench.py --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | tail -1 | \
    python -c "import json,sys,os; d=json.loads(sys.stdin.read()); print('slots', os.environ.get('S'), round(d['value']), round(d['ms_per_step'],1), d['host_profile_last_step'])"
done
