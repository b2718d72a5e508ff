#!/bin/bash
# A/B of a run-time switch on the GPU box: tools/ab_env.sh <workload> <ENVVAR> <value> [<value> ...]
wl=$1; var=$2; shift 2
for round in 1 2; do
for v in "$@"; do
  env $var=$v python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu-baseline --no-extras --sustain 0.3 2>/dev/null \
    | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('$wl', '$var=$v', 'ms', d['ms_per_step'], 'Mrays/s', d['value'], 'e2e', d['e2e']['value'], 'e2e ms', d['e2e']['ms_per_step'])"
done
done
