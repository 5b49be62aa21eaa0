timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 200 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 600 python bench.py --steps 20 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_line.json
