# two-rank sanity run of the default bench line (what the driver's scaling run launches at N = 2)
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/two_rank.json 2> gpurun_out/two_rank.err; tail -2 gpurun_out/two_rank.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/two_rank.json').read().strip().splitlines()[-1])
print('n_gpus',d['n_gpus'],'value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'replay',round(d['replay']['value'],1),'svn',round(d['svn']['value'],2),d['svn'].get('sharded_equals_single_gpu'),'clocks',d['clocks'])
PY
