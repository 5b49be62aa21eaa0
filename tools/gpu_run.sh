mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r1n_pytest.log 2>&1; tail -3 gpurun_out/r1n_pytest.log
python tools/prof_step.py 3 > gpurun_out/r1n_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1n_launches.csv python tools/prof_step.py 2 > gpurun_out/r1n_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'voxel_finalize|voxel_sum' -s 3 -c 4 -o gpurun_out/r1n_build python tools/prof_step.py 2 > gpurun_out/r1n_ncu2.log 2>&1
cat gpurun_out/r1n_plain.log
