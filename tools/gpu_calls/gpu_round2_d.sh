#!/bin/bash
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 ) > gpurun_out/r2d_pytest.log 2>&1
timeout 300 python tools/quick_build.py > gpurun_out/r2d_build.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2d_build_launches.csv python tools/quick_build.py --reps 1 > /dev/null 2>&1
cat gpurun_out/r2d_pytest.log gpurun_out/r2d_build.log
