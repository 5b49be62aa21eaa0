# round 2: 8-GPU verification: sharded SVN == single GPU at 2/4/8 ranks, SVN strong scaling 1/2/4/8, default line at N=8
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2q_topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/r2q_multi.log 2>&1; tail -8 gpurun_out/r2q_multi.log
for N in 1 2 4 8; do
  if [ $N = 1 ]; then
    timeout 600 python bench.py --gpus 1 --workload svn --steps 8 --warmup 3 > gpurun_out/r2q_svn$N.json 2> gpurun_out/r2q_svn$N.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520+N)) bench.py --gpus $N --workload svn --steps 8 --warmup 3 > gpurun_out/r2q_svn$N.json 2> gpurun_out/r2q_svn$N.err
  fi
  tail -2 gpurun_out/r2q_svn$N.err; cut -c1-600 gpurun_out/r2q_svn$N.json; echo
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29540 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2q_bench8.json 2> gpurun_out/r2q_bench8.err; tail -3 gpurun_out/r2q_bench8.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r2q_bench8.json').read().strip().splitlines()[-1])
    print('value',d['value'],'e2e',d['e2e']['value']); print('replay',d['replay']); print('svn',d['svn'])
except Exception as e: print('no line',e)
PY
