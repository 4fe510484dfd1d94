#!/bin/bash
# tools/ab_env.sh "ENV1=..." "ENV2=..." -- headline device bench under different environment settings
mkdir -p gpurun_out
for e in "$@"; do
  env $e python bench.py --steps 30 --warmup 5 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('$e', round(d['ms_per_step'],4), round(d['value'],1), json.dumps(d['stages_ms_per_step']))" | tee -a gpurun_out/ab_env.log
done
