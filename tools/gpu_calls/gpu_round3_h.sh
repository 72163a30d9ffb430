#!/bin/bash
# round 3, call h: the round-end sequence on the final tree — build check, smoke(), full GPU suite, both bench arms at N=1
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3h_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/r3h_smoke.log
python -m pytest tests -m gpu -x -q > gpurun_out/r3h_gpu_suite.log 2>&1; echo "suite rc=$?"; tail -n 3 gpurun_out/r3h_gpu_suite.log
python bench.py --impl reference > gpurun_out/r3h_bench_reference.json 2> gpurun_out/r3h_bench_reference.err; echo "reference arm rc=$?"
python bench.py > gpurun_out/r3h_bench_n1.json 2> gpurun_out/r3h_bench_n1.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/r3h_bench_n1.json
