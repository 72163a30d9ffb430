#!/bin/bash
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 ) > gpurun_out/r2c_pytest.log 2>&1
timeout 300 python tools/quick_trace.py --reps 5 > gpurun_out/r2c_trace.log 2>&1
timeout 300 python tools/quick_illum.py --reps 2 >> gpurun_out/r2c_trace.log 2>&1
timeout 300 python tools/quick_build.py > gpurun_out/r2c_build.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2c_build_launches.csv python tools/quick_build.py --reps 1 > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_collapse -s 1 -c 1 -f -o gpurun_out/r2c_collapse python tools/quick_build.py --reps 1 > /dev/null 2>&1
ncu -i gpurun_out/r2c_collapse.ncu-rep --page raw --csv > gpurun_out/r2c_collapse_raw.csv 2>/dev/null
ncu -i gpurun_out/r2c_collapse.ncu-rep --page source --csv > gpurun_out/r2c_collapse_src.csv 2>/dev/null
rm -f gpurun_out/r2c_collapse.ncu-rep
cat gpurun_out/r2c_pytest.log gpurun_out/r2c_trace.log gpurun_out/r2c_build.log
