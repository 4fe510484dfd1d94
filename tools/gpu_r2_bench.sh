#!/bin/bash
# tools/gpu_r2_bench.sh N -- the driver's bench command at N GPUs (ours, then the reference arm), lines kept in gpurun_out/
N=${1:-1}; mkdir -p gpurun_out
if [ "$N" = 1 ]; then
  RUN="python"
else
  RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
fi
timeout 900 $RUN bench.py --gpus $N --steps 30 --warmup 5 2> gpurun_out/bench_ours_n$N.err | grep '^{' > gpurun_out/bench_ours_n$N.json; echo "ours rc=$?"
tail -n 5 gpurun_out/bench_ours_n$N.err
if [ "$2" = ref ]; then
  timeout 900 $RUN bench.py --impl reference --gpus $N --steps 30 --warmup 5 2> gpurun_out/bench_ref_n$N.err | grep '^{' > gpurun_out/bench_ref_n$N.json; echo "ref rc=$?"
fi
python - <<PY
import json
for who in ("ours", "ref"):
    try: d = json.load(open("gpurun_out/bench_%s_n$N.json" % who))
    except Exception as e: print(who, "no line", e); continue
    print(who, "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 4), "e2e", d.get("e2e"))
    print("   stages", d.get("stages_ms_per_step")); print("   exchange", d.get("exchange")); print("   window_c4", d.get("window_c4"))
PY
