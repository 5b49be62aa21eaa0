timeout 600 python bench.py --workload replay --steps 5 2>&1 | tail -1 | tee gpurun_out/bench_replay_1gpu.json | cut -c1-1800
timeout 900 python -m pytest tests/test_gpu_large_map.py -x -q -k replay 2>&1 | tail -3
