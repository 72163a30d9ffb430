#!/bin/bash
mkdir -p gpurun_out
( timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 ) > gpurun_out/r2e_pytest.log 2>&1
( time timeout 1500 python bench.py > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err ) 2> gpurun_out/r2e_bench.time
tail -5 gpurun_out/r2e_bench.err
cat gpurun_out/r2e_pytest.log gpurun_out/r2e_bench.time
