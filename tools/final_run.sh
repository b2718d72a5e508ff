#!/bin/bash
# tools/final_run.sh <outdir>: the round's verification on one B200: GPU test-suite, build comparison, both bench arms, ncu
# launch list and one --set full capture of the dominant kernel (each ncu pass only after the plain run exited 0).
o=$1; mkdir -p $o
(python -m pytest tests -m gpu -q -s 2>&1 | tail -400) > $o/pytest_gpu.log; tail -3 $o/pytest_gpu.log
python tools/bench_build.py > $o/build_device_vs_host.jsonl 2> $o/build.err; cat $o/build_device_vs_host.jsonl | cut -c1-600
python bench.py > $o/bench.json 2> $o/bench.err && echo bench ok
python bench.py --impl reference --steps 3 --warmup 1 > $o/bench_reference_arm.json 2> $o/bench_ref.err
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > /dev/null 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > $o/ncu1.log 2>&1
SNL_RENDER_VARIANT=1 ncu --set full --clock-control none --import-source on -k regex:k_render --launch-skip 3 --launch-count 1 -o $o/k_render -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras --sustain 0.05 > $o/ncu2.log 2>&1
ls -la $o
