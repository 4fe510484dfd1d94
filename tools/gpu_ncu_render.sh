#!/bin/bash
# tools/gpu_ncu_render.sh v1 v2 ... -- ncu --set full of the two render kernels for library variants ("base" = product)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-window-c4"
for v in "$@"; do
  if [ $v = base ]; then unset GSR_B200_LIB; else export GSR_B200_LIB=$PWD/monogs_b200/libgsr_b200_$v.so; fi
  $CMD > gpurun_out/plain_$v.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:render_ -s 2 -c 4 -f -o gpurun_out/prof_render_$v $CMD > gpurun_out/ncu_full_$v.log 2>&1
  echo "$v ncu rc=$?"
done
