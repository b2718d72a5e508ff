#!/bin/bash
# A/B of libsnail_b200 tuning variants on the GPU box: tools/ab.sh <workload> <suffix> [<suffix> ...]   ("-" = default lib)
# Prints ms/step and Mrays/s of bench.py (no CPU baseline, no extras) for each variant, two rounds.
wl=$1; shift
for round in 1 2; do
for v in "$@"; do
  sfx=$v; [ "$v" = "-" ] && sfx=""
  SNL_LIB=$PWD/snail_b200/libsnail_b200$sfx.so python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu-baseline --no-extras --sustain 0.3 2>/dev/null \
    | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('$wl', '${v}', 'ms', d['ms_per_step'], 'Mrays/s', d['value'], 'sustained', d.get('sustained',{}).get('value'))"
done
done
