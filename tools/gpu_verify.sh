mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/v_pytest.log 2>&1; tail -5 gpurun_out/v_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/v_smoke.log 2>&1; tail -3 gpurun_out/v_smoke.log
timeout 600 python bench.py > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err; tail -2 gpurun_out/v_bench.err; cat gpurun_out/v_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/v_bench_ref.json 2> gpurun_out/v_bench_ref.err; cat gpurun_out/v_bench_ref.json
timeout 600 python bench.py --workload multigrid > gpurun_out/v_bench_mg.json 2> gpurun_out/v_bench_mg.err; cat gpurun_out/v_bench_mg.json
