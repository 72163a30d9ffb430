#!/bin/bash
# round-2 GPU call L: ballot ranking in the sort (parity + timings), CUDA-line profiles of k_collapse and k_sort_pass
mkdir -p gpurun_out
O=gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_build.py tests/test_gpu_trace.py tests/test_gpu_sss.py -x -q 2>&1 | tail -8 ) > $O/r2l_pytest.log 2>&1
timeout 300 python tools/quick_build.py --reps 3 > $O/r2l_build.log 2>&1
timeout 300 python tools/quick_build.py --plants 500 --reps 2 > $O/r2l_build_c3.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_collapse|k_sort_pass' -s 12 -c 4 -f -o $O/r2l_bk \
    python tools/quick_build.py --reps 1 > $O/r2l_bk_ncu.log 2>&1
ncu -i $O/r2l_bk.ncu-rep --page source --print-source cuda --csv > $O/r2l_bk_cuda.csv 2> $O/r2l_bk_cuda.err
ncu -i $O/r2l_bk.ncu-rep --page raw --csv > $O/r2l_bk_raw.csv
rm -f $O/r2l_bk.ncu-rep
cat $O/r2l_pytest.log $O/r2l_build.log $O/r2l_build_c3.log; tail -n 3 $O/r2l_bk_ncu.log; wc -l $O/r2l_bk_cuda.csv
