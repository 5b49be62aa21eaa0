mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r02v_bench.json 2> gpurun_out/r02v_bench.err; tail -3 gpurun_out/r02v_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02v_ref.json 2> gpurun_out/r02v_ref.err; tail -2 gpurun_out/r02v_ref.err; cut -c1-300 gpurun_out/r02v_ref.json
timeout 600 python bench.py --workload multigrid --no-cpu-baseline > gpurun_out/r02v_mg.json 2> gpurun_out/r02v_mg.err; tail -2 gpurun_out/r02v_mg.err; cut -c1-200 gpurun_out/r02v_mg.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02v_bench.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'],'serial',d['e2e_serial']['value'],'build_ms',d['config']['build_ms'],'align_ms',d['config']['align_ms'])
print('roofline',d['roofline']['frac'],d['roofline']['kernel_us'],d['roofline'].get('session_eval_us'),d['roofline'].get('issue'))
print('roofline_build',d['roofline_build']['frac'],d['roofline_build']['build_us']); print('parity',d['parity']['ok']); print('replay',d['replay']['value']); print('svn',d['svn']['value']); print('cpu',d['cpu_baseline']['value'])
PY
