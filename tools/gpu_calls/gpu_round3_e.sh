#!/bin/bash
# round 3, call e: trace-side estimate of the two-level (instance) BVH on the C2 / C4 / C3 scenes; GPU suite incl. the T20 fixture test
mkdir -p gpurun_out
for p in 50 250 500; do
  python tools/exp_two_level.py --plants $p > gpurun_out/r3e_two_level_$p.log 2>&1; echo "two-level $p rc=$?"; cat gpurun_out/r3e_two_level_$p.log | tail -n 8
done
python -m pytest tests -m gpu -x -q > gpurun_out/r3e_gpu_suite.log 2>&1; echo "suite rc=$?"; tail -n 4 gpurun_out/r3e_gpu_suite.log
