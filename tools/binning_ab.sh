#!/bin/bash
# tools/binning_ab.sh [workloads] -- stage times of the three binning implementations (GSR_BINNING = two-level default / single / radix)
mkdir -p gpurun_out
for w in ${1:-headline c3}; do
for m in twolevel single radix; do
  GSR_BINNING=$m python bench.py --workload $w --steps 30 --warmup 5 --no-e2e --no-cpu-baseline --no-window-c4 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('$w', '$m', round(d['ms_per_step'],4), json.dumps(d['stages_ms_per_step']))" | tee -a gpurun_out/binning_ab.log
done
done
