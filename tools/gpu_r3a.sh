mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigrid.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python bench.py --workload multigrid --no-cpu-baseline > gpurun_out/r3a_mg.json 2> gpurun_out/r3a_mg.err; tail -2 gpurun_out/r3a_mg.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3a_mg.json').read().strip().splitlines()[-1])
print('mg value',d['value'],d['ms_per_step'], {k:v for k,v in d['config'].items() if 'us' in k or 'ms' in k})
print(d.get('roofline'))
PY
