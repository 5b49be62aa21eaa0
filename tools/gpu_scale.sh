# usage: bash tools/gpu_scale.sh N   (run under gpurun --gpus N)
N=$1
mkdir -p gpurun_out
for w in svn replay map_align; do
  if [ "$N" = "1" ]; then
    timeout 600 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/scale_${N}_$w.json 2> gpurun_out/scale_${N}_$w.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --workload $w --steps 10 --warmup 3 > gpurun_out/scale_${N}_$w.json 2> gpurun_out/scale_${N}_$w.err
  fi
  python -c "
import json; d=json.loads(open('gpurun_out/scale_${N}_$w.json').read().strip().splitlines()[-1]); print('$w', d['n_gpus'], d['value'], d['unit'], d['ms_per_step'], d.get('e2e',{}).get('value'), d['scaling'])" || tail -5 gpurun_out/scale_${N}_$w.err
done
