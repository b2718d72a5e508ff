#!/bin/bash
# A/B of the host-facing render path on the GPU box: tools/ab_e2e.sh <workload>   (row bands on / off)
wl=$1
for round in 1 2; do
for b in 1 0; do
  SNL_RENDER_BANDS=$b python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu-baseline --no-extras --sustain 0.3 2>/dev/null \
    | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('$wl', 'bands=$b', 'ms', d['ms_per_step'], 'Mrays/s', d['value'], 'e2e', d['e2e']['value'], 'e2e ms', d['e2e']['ms_per_step'])"
done
done
