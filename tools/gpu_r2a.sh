# round 2, first verification on one B200: new parity tests first (fail fast), then the whole GPU suite, smoke, bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_baseline_configs.py tests/test_gpu_svn.py -m gpu -x -q > gpurun_out/r2a_new.log 2>&1; tail -15 gpurun_out/r2a_new.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2a_pytest.log 2>&1; tail -15 gpurun_out/r2a_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1; tail -3 gpurun_out/r2a_smoke.log
timeout 900 python bench.py > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; tail -5 gpurun_out/r2a_bench.err; cut -c1-600 gpurun_out/r2a_bench.json
