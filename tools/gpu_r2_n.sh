#!/bin/bash
# tools/gpu_r2_n.sh N -- multi-GPU round: exchange test + check (phase trace), then the bench at N GPUs
N=${1:-2}; mkdir -p gpurun_out
python -m pytest tests/test_window_exchange.py -m gpu -q --tb=short 2>&1 | tail -5
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/xchg_check.py --keyframes 11 --out gpurun_out/xchg_check_n$N.json > gpurun_out/xchg_check_n$N.log 2>&1; echo "xchg rc=$?"
tail -3 gpurun_out/xchg_check_n$N.log | cut -c1-1500
bash tools/gpu_r2_bench.sh $N $2
