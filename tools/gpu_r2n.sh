mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r2n_bench.json 2> gpurun_out/r2n_bench.err; tail -3 gpurun_out/r2n_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2n_bench.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],'build_ms',d['config']['build_ms'],'align_ms',d['config']['align_ms'])
print('roofline',d['roofline']['frac'],d['roofline']['kernel_us'],d['roofline'].get('session_eval_us'))
print('roofline_build',d['roofline_build']); print('parity',d['parity']['ok']); print('replay',d['replay']['value']); print('svn',d['svn']['value'])
PY
