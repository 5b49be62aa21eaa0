mkdir -p gpurun_out
for k in 2 3 1; do
  NDTB_BATCH_IN_FLIGHT=$k timeout 600 python bench.py --workload replay --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/replay_$k.json 2> gpurun_out/replay_$k.err; tail -1 gpurun_out/replay_$k.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/replay_$k.json').read().strip().splitlines()[-1])
print('in flight $k:', round(d['value'],1), 'scans/s', {kk:vv for kk,vv in d['config'].items() if 'us' in kk or 'evaluations' in kk})
PY
done
timeout 600 python -m pytest tests/test_gpu_large_map.py -m gpu -x -q -k replay 2>&1 | tail -3
