#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x --tb=short > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 30 gpurun_out/pytest_gpu.log
rm -f gpurun_out/variants.log
bash tools/variant_bench.sh -w "headline c3" "$@"
