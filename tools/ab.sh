#!/bin/bash
# A/B of library variants on the GPU box: bash tools/ab.sh name1[:ENV=V] name2 ...   (variants in tools/scratch/var/<name>.so)
mkdir -p gpurun_out
for spec in "$@"; do
  name=${spec%%:*}; envs=""
  [ "$spec" != "$name" ] && envs=${spec#*:}
  cp tools/scratch/var/$name.so my_orb_slam2_b200/liborbfe.so
  for rep in 1; do
    env $envs python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-matchers --no-shapes --latency-frames 100 2>gpurun_out/ab_$name.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$spec', 'value %.0f e2e %.0f p50 %.4f' % (d['value'], d['e2e']['value'], d['latency_ms_p50']), {k: round(v,4) for k,v in d['stages_ms_per_launch'].items()})
"
  done
done
