#!/bin/bash
# Time bench.py with each library variant under asimov-mjlab_b200/_variants (experiment helper; run on the GPU box).
for v in asimov-mjlab_b200/_variants/*.so; do
  cp $v asimov-mjlab_b200/libb200sim.so
  python bench.py --steps ${STEPS:-150} --warmup 10 --cpu-baseline 0 ${BENCH_ARGS} 2>&1 | tail -1 | python -c "
import sys, json
j = json.loads(sys.stdin.read())
print('$(basename $v)', round(j['value'] / 1e6, 3), 'M/s  solve', round(j['roofline']['stage_ms']['solve'], 4), 'ms  step', round(j['ms_per_step'], 3))"
done
