#!/bin/bash
# tools/scale_one.sh <outdir> <tag> <N> <workload> <res> [bench args...]: one bench.py run under torchrun, one summary line
out=$1; tag=$2; n=$3; wl=$4; res=$5; shift 5
mkdir -p $out
port=$((29700 + RANDOM % 200))
if [ "$n" = 1 ]; then
  python bench.py --gpus 1 --workload $wl --res $res --no-cpu-baseline --no-extras "$@" 2> $out/err_$tag.log
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --workload $wl --res $res --no-cpu-baseline --no-extras "$@" 2> $out/err_$tag.log
fi | tail -1 | tee -a $out/scale.jsonl | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('$tag', 'N', d['n_gpus'], 'value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'e2e ms', d['e2e']['ms_per_step'], 'sustained', d['sustained']['value'])" || { echo "$tag FAILED"; tail -5 $out/err_$tag.log; }
