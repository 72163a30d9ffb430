#!/bin/bash
# round 3, call b (2 GPUs): the multi-GPU tests, incl. the agreement step before the collectives (one rank failing)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/r3b_multi.log 2>&1; echo "multi rc=$?" | tee -a gpurun_out/r3b_multi.log
tail -n 30 gpurun_out/r3b_multi.log
