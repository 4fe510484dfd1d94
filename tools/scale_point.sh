# one point of the scaling table: bash tools/scale_point.sh N [workload]
N=$1; WL=${2:-headline}
mkdir -p gpurun_out
if [ "$N" = 1 ]; then
  timeout 600 python bench.py --gpus 1 --steps 30 --warmup 5 --workload $WL > gpurun_out/scale_${WL}_n$N.json 2> gpurun_out/scale_${WL}_n$N.err
else
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 30 --warmup 5 --workload $WL 2> gpurun_out/scale_${WL}_n$N.err | grep '^{' > gpurun_out/scale_${WL}_n$N.json
fi
python -c "
import json,sys
d=json.load(open('gpurun_out/scale_${WL}_n$N.json'))
print('$WL N=$N', round(d['value'],1), 'it/s', round(d['ms_per_step'],4), 'ms', d.get('exchange',{}).get('mode'), d.get('exchange',{}).get('times_ms'))"
