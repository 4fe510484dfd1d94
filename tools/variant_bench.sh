#!/bin/bash
# tools/variant_bench.sh [-w "headline c2"] v1 v2 ... -- bench stage times of experimental library variants
# (monogs_b200/libgsr_b200_<v>.so, built with `python -m monogs_b200.build --variant <v> DEFINE=..`)
mkdir -p gpurun_out
WL="headline"
if [ "$1" = "-w" ]; then WL="$2"; shift 2; fi
for w in $WL; do
for v in base "$@"; do
  if [ $v = base ]; then unset GSR_B200_LIB; else export GSR_B200_LIB=$PWD/monogs_b200/libgsr_b200_$v.so; fi
  python bench.py --workload $w --steps 30 --warmup 5 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('$w', '$v', round(d['ms_per_step'],4), json.dumps(d['stages_ms_per_step']))" | tee -a gpurun_out/variants.log
done
done
