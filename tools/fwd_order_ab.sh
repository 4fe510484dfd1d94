mkdir -p gpurun_out
for w in headline c2 c3; do
for o in 0 1; do
GSR_FWD_TILE_ORDER=$o python bench.py --workload $w --steps 30 --warmup 5 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read())
s=d['stages_ms_per_step']
print('$w fwd_order=$o', round(d['ms_per_step'],4), 'render_fwd', s['render_fwd'], 'render_bwd', s['render_bwd'])" | tee -a gpurun_out/fwd_order_ab.log
done
done
GSR_FWD_TILE_ORDER=1 python -m pytest tests/test_gpu_parity.py tests/test_gpu_api.py -m gpu -x -q 2>&1 | tail -2
