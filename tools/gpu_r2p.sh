mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_svn.py tests/test_gpu_multi.py tests/test_gpu_baseline_configs.py -m gpu -x -q 2>&1 | tail -8
timeout 600 python bench.py --gpus 1 --workload svn --steps 8 --warmup 3 > gpurun_out/r2p_svn1.json 2> gpurun_out/r2p_svn1.err; tail -2 gpurun_out/r2p_svn1.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --workload svn --steps 8 --warmup 3 > gpurun_out/r2p_svn2.json 2> gpurun_out/r2p_svn2.err; tail -2 gpurun_out/r2p_svn2.err
python - <<'PY'
import json
for n in (1,2):
    try:
        d=json.loads(open(f'gpurun_out/r2p_svn{n}.json').read().strip().splitlines()[-1])
        print(n,'value',d['value'],{k:d['config'].get(k) for k in ('eval_us_per_iter','stein_us_per_iter','allreduce_us_per_iter','iterations','sharded_equals_single_gpu')})
    except Exception as e: print(n,'no line',e)
PY
