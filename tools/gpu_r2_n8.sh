#!/bin/bash
# tools/gpu_r2_n8.sh N -- exchange parity + phases, a LOW-OCCLUSION exchange measurement, then the bench at N GPUs
N=${1:-8}; mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $T --master-port 29511 tools/xchg_check.py --keyframes 11 --out gpurun_out/xchg_check_n$N.json > gpurun_out/xchg_check_n$N.log 2>&1; echo "xchg rc=$?"
timeout 300 $T --master-port 29513 tools/xchg_check.py --kind replica --gaussians 200000 --keyframes $N --opacity-scale 0.08 --out gpurun_out/xchg_lowocc_n$N.json > gpurun_out/xchg_lowocc_n$N.log 2>&1; echo "lowocc rc=$?"
python - <<PY
import json
for f in ("xchg_check_n$N", "xchg_lowocc_n$N"):
    try: d = json.load(open("gpurun_out/%s.json" % f))
    except Exception as e: print(f, "no report", e); continue
    print(f, "pass", d["pass"], "worst", d["worst_rel_err_all_ranks"], "touched_union", d.get("touched_fraction_union"))
    print("   phases_us", d.get("sparse_phases_us")); print("   report", d.get("sparse_report")); print("   grow", d.get("grow"))
PY
bash tools/gpu_r2_bench.sh $N $2
