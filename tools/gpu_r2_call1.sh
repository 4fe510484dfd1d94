#!/bin/bash
# round 2, call 1: all GPU tests (no -x: full picture), smoke, bench line of the round-1 kernels on today's box,
# compute-sanitizer memcheck over the small parity cases.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -rA --tb=short --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
python bench.py > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench rc=$?" >> gpurun_out/bench_ours.err
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest tests/test_gpu_parity.py tests/test_gpu_api.py -m gpu -q -x \
  -k "golden or edge or cpu_oracle or mark_visible or single_call or rows_backward or knn" > gpurun_out/sanitizer_memcheck.log 2>&1
echo "memcheck rc=$?" >> gpurun_out/sanitizer_memcheck.log
tail -n 25 gpurun_out/pytest_gpu.log; tail -n 3 gpurun_out/smoke.log gpurun_out/bench_ours.err; tail -n 12 gpurun_out/sanitizer_memcheck.log
cut -c1-1200 gpurun_out/bench_ours.json
