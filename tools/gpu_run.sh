mkdir -p gpurun_out
python bench.py --workload svn --steps 3 --warmup 1 > gpurun_out/r1t_svn1.json 2> gpurun_out/r1t_svn1.err; cat gpurun_out/r1t_svn1.json; tail -3 gpurun_out/r1t_svn1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload svn --steps 3 --warmup 1 > gpurun_out/r1t_svn2.json 2> gpurun_out/r1t_svn2.err; cat gpurun_out/r1t_svn2.json; tail -3 gpurun_out/r1t_svn2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r1t_map2.json 2> gpurun_out/r1t_map2.err; cat gpurun_out/r1t_map2.json; tail -3 gpurun_out/r1t_map2.err
