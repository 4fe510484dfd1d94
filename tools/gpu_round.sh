#!/bin/bash
# tools/gpu_round.sh -- what one gpurun call of a build->measure cycle runs on the B200 box.
# usage: bash tools/gpu_round.sh [tests] [bench] [ref] [launches] [ncufull]
mkdir -p gpurun_out
for what in "$@"; do
case $what in
tests)
  python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
  python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log ;;
bench)
  python bench.py > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench rc=$?" >> gpurun_out/bench_ours.err ;;
ref)
  python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?" >> gpurun_out/bench_ref.err ;;
launches)
  CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-window-c4"
  $CMD > gpurun_out/plain.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1 ;;
ncufull)
  CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-window-c4"
  $CMD > gpurun_out/plain2.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:render_ -s 2 -c 4 -f -o gpurun_out/prof_render $CMD > gpurun_out/ncu_full.log 2>&1 ;;
ncuother)
  CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-window-c4"
  $CMD > gpurun_out/plain3.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:"onesweep|preprocess_|gather_|tile_|chunk_|tau_|digit_|refine_" -s 21 -c 21 -f -o gpurun_out/prof_other $CMD > gpurun_out/ncu_other.log 2>&1 ;;
ncupre)
  CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-window-c4"
  $CMD > gpurun_out/plain5.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:"preprocess_|gather_rects|chunk_scan" -s 4 -c 4 -f -o gpurun_out/prof_pre $CMD > gpurun_out/ncu_pre.log 2>&1 ;;
ncuts)
  CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-window-c4"
  $CMD > gpurun_out/plain4.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:"tile_count|tile_scatter|tile_offsets" -s 3 -c 3 -f -o gpurun_out/prof_ts $CMD > gpurun_out/ncu_ts.log 2>&1 ;;
esac
done
tail -n 3 gpurun_out/pytest_gpu.log gpurun_out/smoke.log gpurun_out/bench_ours.err gpurun_out/bench_ref.err 2>/dev/null
cat gpurun_out/bench_ours.json gpurun_out/bench_ref.json 2>/dev/null | cut -c1-1500
