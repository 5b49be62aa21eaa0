mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1s_bench_small.json 2> gpurun_out/r1s_bench_small.err &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1s_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1s_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:eval_kernel_pclomp -s 60 -c 3 -o gpurun_out/r1s_eval python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1s_ncu2.log 2>&1
cat gpurun_out/r1s_bench_small.json; tail -2 gpurun_out/r1s_bench_small.err
python bench.py --steps 20 --warmup 3 > gpurun_out/r1s_bench.json 2> gpurun_out/r1s_bench.err; cat gpurun_out/r1s_bench.json; tail -3 gpurun_out/r1s_bench.err
