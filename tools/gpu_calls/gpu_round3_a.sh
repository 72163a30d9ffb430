#!/bin/bash
# round 3, call a: the no-allocation flux entry point alone in a fresh process, then the whole GPU suite (it runs last there)
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r3a_noalloc_alone.log 2>&1; echo "alone rc=$?" | tee -a gpurun_out/r3a_noalloc_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/r3a_gpu_suite.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/r3a_gpu_suite.log
tail -5 gpurun_out/r3a_noalloc_alone.log gpurun_out/r3a_gpu_suite.log
