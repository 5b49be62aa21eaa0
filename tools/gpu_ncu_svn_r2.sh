mkdir -p gpurun_out
python bench.py --workload svn --steps 2 --warmup 1 > gpurun_out/r2t_plain.json 2> gpurun_out/r2t_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:'stein_kernel|eval_kernel' -s 12 -c 4 -f -o gpurun_out/r2t_svn python bench.py --workload svn --steps 2 --warmup 1 > gpurun_out/r2t_ncu.log 2>&1
tail -2 gpurun_out/r2t_ncu.log | cut -c1-200
