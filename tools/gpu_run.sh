timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/bench_identity.json 2> gpurun_out/bench_identity.err; tail -c 3000 gpurun_out/bench_identity.json
