#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/exp_illum.py > gpurun_out/r2f_exp_illum.log 2>&1
( timeout 600 python -m pytest tests/test_gpu_entrypoints.py -x -q 2>&1 | tail -8 ) > gpurun_out/r2f_pytest.log 2>&1
timeout 600 python bench.py --configs "" > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err
cat gpurun_out/r2f_exp_illum.log gpurun_out/r2f_pytest.log; tail -3 gpurun_out/r2f_bench.err
