python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload replay --steps 5 2>&1 | tail -1 > gpurun_out/bench_replay_2gpu.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 > gpurun_out/bench_map_align_2gpu.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --workload svn --steps 5 2>&1 | tail -1 > gpurun_out/bench_svn_2gpu.json
for f in replay map_align svn; do python -c "import json,sys; d=json.load(open('gpurun_out/bench_${f}_2gpu.json')); print('$f', d['n_gpus'], d['value'], d['e2e']['value'])"; done
