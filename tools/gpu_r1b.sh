set -x
mkdir -p gpurun_out
python tools/prof_step.py 2 > gpurun_out/r1b_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1b_launches.csv python tools/prof_step.py 2 > gpurun_out/r1b_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:eval_kernel -s 21 -c 2 -o gpurun_out/r1b_eval python tools/prof_step.py 2 > gpurun_out/r1b_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'^(?!.*eval_kernel)' -c 40 -o gpurun_out/r1b_build python tools/prof_step.py 1 > gpurun_out/r1b_ncu3.log 2>&1
cat gpurun_out/r1b_plain.log
ls -la gpurun_out
