# round 2: multi-GPU SVN (2 GPUs): peer-written records, sharded == single GPU; new svn KDTREE / calculateScore / export tests; bench at N=2
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2b_topo.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/r2b_multi.log 2>&1; tail -15 gpurun_out/r2b_multi.log
timeout 900 python -m pytest tests/test_gpu_svn.py tests/test_gpu_pclomp.py tests/test_facade_cpp.py -m gpu -q > gpurun_out/r2b_tests.log 2>&1; tail -15 gpurun_out/r2b_tests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2b_bench2.json 2> gpurun_out/r2b_bench2.err; tail -5 gpurun_out/r2b_bench2.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r2b_bench2.json').read().strip().splitlines()[-1])
    print('value',d['value'],'e2e',d['e2e']['value']); print('replay',d['replay']); print('svn',d['svn'])
except Exception as e: print('no line',e)
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload svn --steps 10 --warmup 3 > gpurun_out/r2b_svn2.json 2> gpurun_out/r2b_svn2.err; tail -3 gpurun_out/r2b_svn2.err; cut -c1-400 gpurun_out/r2b_svn2.json
