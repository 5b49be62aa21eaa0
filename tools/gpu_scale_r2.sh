# final round-2 multi-GPU verification on one 8 x B200 box
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/r02s_multi.log 2>&1; tail -3 gpurun_out/r02s_multi.log
for N in 1 2 4 8; do
  if [ $N = 1 ]; then
    timeout 600 python bench.py --gpus 1 --workload svn --steps 8 --warmup 3 > gpurun_out/r02s_svn$N.json 2> gpurun_out/r02s_svn$N.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29620+N)) bench.py --gpus $N --workload svn --steps 8 --warmup 3 > gpurun_out/r02s_svn$N.json 2> gpurun_out/r02s_svn$N.err
  fi
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r02s_svn$N.json').read().strip().splitlines()[-1]); s=d['svn']
    print($N,'svn',round(d['value'],2),'ms',round(d['ms_per_step'],3),'eval',round(s['eval_us_per_iter'],1),'stein',round(s['stein_us_per_iter'],1),'wait',round(s['allreduce_us_per_iter'],1),s['sharded_equals_single_gpu'])
except Exception as e: print($N,'no line',e)
PY
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29640 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02s_bench8.json 2> gpurun_out/r02s_bench8.err; tail -2 gpurun_out/r02s_bench8.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r02s_bench8.json').read().strip().splitlines()[-1])
    print('N=8 value',d['value'],'e2e',d['e2e']['value'],'serial',d['e2e_serial']['value']); print('replay',d['replay']['value']); print('svn',d['svn']['value'])
except Exception as e: print('no line',e)
PY
