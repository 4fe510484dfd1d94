# 8-GPU confirmation of the sparse window exchange: correctness vs NCCL, then the headline and C4 bench lines.
set -x
mkdir -p gpurun_out
N=${1:-8}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/xchg_check.py --keyframes 11 --out gpurun_out/xchg_check_n$N.json > gpurun_out/xchg_check_n$N.log 2>&1; echo "xchg rc=$?"
tail -2 gpurun_out/xchg_check_n$N.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_n${N}_sparse.log 2>&1; echo "bench rc=$?"
tail -1 gpurun_out/bench_n${N}_sparse.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 30 --warmup 5 --workload c4 > gpurun_out/bench_c4_n${N}_sparse.log 2>&1; echo "bench c4 rc=$?"
tail -1 gpurun_out/bench_c4_n${N}_sparse.log
