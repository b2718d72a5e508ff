#!/bin/bash
# tools/scale_run.sh <outdir> <N...>: bench.py on N GPUs of this box for the headline and the two 4K configs (p2p compose),
# plus the NCCL-gather compose of the headline. One JSON line per run in <outdir>/scale.jsonl.
out=$1; shift
mkdir -p $out
port=29600
run() { # n workload res compose tag
  port=$((port+1))
  if [ "$1" = 1 ]; then
    python bench.py --gpus 1 --workload $2 --res $3 --no-cpu-baseline --no-extras --steps 20 --warmup 3 2> $out/err_$5.log
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $port bench.py --gpus $1 --workload $2 --res $3 --compose $4 --no-cpu-baseline --no-extras --steps 20 --warmup 3 2> $out/err_$5.log
  fi | tail -1 | tee -a $out/scale.jsonl | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('$5', 'N', d['n_gpus'], 'value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e']['value'], 'e2e ms', d['e2e']['ms_per_step'], 'sustained', d['sustained']['value'])" || { echo "$5 FAILED"; tail -5 $out/err_$5.log; }
}
for n in "$@"; do
  run $n sponza_like_whitted 1920x1080 p2p sponza_p2p_n$n
  [ $n -gt 1 ] && run $n sponza_like_whitted 1920x1080 nccl sponza_nccl_n$n
  run $n barracks_shadow 3840x2160 p2p barracks4k_p2p_n$n
  run $n lancia_whitted 3840x2160 p2p lancia4k_p2p_n$n
done
