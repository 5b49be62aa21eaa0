mkdir -p gpurun_out
python bench.py --workload multigrid --steps 3 --warmup 3 > gpurun_out/mg_plain.json 2> gpurun_out/mg_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/mg_launches.csv python bench.py --workload multigrid --steps 3 --warmup 3 > gpurun_out/mg_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:mg_eval_kernel -s 20 -c 1 -o gpurun_out/mg_eval python bench.py --workload multigrid --steps 3 --warmup 3 > gpurun_out/mg_ncu2.log 2>&1
ls -la gpurun_out
