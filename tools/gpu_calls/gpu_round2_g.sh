#!/bin/bash
# round-2 GPU call G (session 4 baseline): full GPU suite, full bench, launch list, full ncu capture of k_trace_rays
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/r2g_pytest.log 2>&1
( time timeout 1500 python bench.py > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err ) 2> gpurun_out/r2g_bench.time
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2g_bench_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r2g_bench_ncu.log 2>&1
timeout 900 bash tools/ncu_trace.sh r2g > gpurun_out/r2g_ncu_trace.log 2>&1
rm -f gpurun_out/r2g_trace.ncu-rep
cat gpurun_out/r2g_pytest.log gpurun_out/r2g_bench.time; tail -3 gpurun_out/r2g_bench.err; tail -30 gpurun_out/r2g_ncu_trace.log
