mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for w in headline c2 c3 c5; do
for o in 0 1; do
GSR_TILE_ORDER=$o python bench.py --workload $w --steps 30 --warmup 5 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read())
print('$w order=$o', round(d['ms_per_step'],4), json.dumps(d['stages_ms_per_step']))" | tee -a gpurun_out/tile_order_ab.log
done
done
