#!/bin/bash
# tools/bench_configs.sh -- the non-headline BASELINE.json configs, ours and the reference arm, one JSON line each
mkdir -p gpurun_out
: > gpurun_out/bench_configs.jsonl
for wl in c2 c3 c4 c5; do
  for impl in ours reference; do
    python bench.py --workload $wl --impl $impl --steps 20 --warmup 5 --no-cpu-baseline 2>> gpurun_out/bench_configs.err >> gpurun_out/bench_configs.jsonl
  done
done
python - <<'PY'
import json
for l in open('gpurun_out/bench_configs.jsonl'):
    d = json.loads(l)
    print(d['config']['workload'], d.get('impl', 'ours'), round(d['value'], 1), 'it/s', round(d['ms_per_step'], 3), 'ms',
          'e2e', round(d['e2e']['value'], 1) if d.get('e2e') else None, d['config'].get('units'))
PY
