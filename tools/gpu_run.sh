timeout 900 python -m pytest tests/test_multigrid.py -m gpu -q 2>&1 | tail -12
python bench.py --workload multigrid > gpurun_out/bench_mg.json 2> gpurun_out/bench_mg.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_mg.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['roofline']['kernel_us'], d['config']['evaluations_per_align'])
PY
tail -3 gpurun_out/bench_mg.err
