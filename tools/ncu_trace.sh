#!/bin/bash
# ncu capture of k_trace_rays (after the same command exited 0 without ncu).  usage: tools/ncu_trace.sh <tag>
set -e
tag=${1:-cur}
python tools/quick_trace.py --reps 2 --check 0 > gpurun_out/${tag}_plain.log 2>&1
cat gpurun_out/${tag}_plain.log
ncu --set full --clock-control none --import-source on -k regex:k_trace_rays -s 3 -c 1 -f -o gpurun_out/${tag}_trace \
    python tools/quick_trace.py --reps 2 --check 0 > gpurun_out/${tag}_ncu.log 2>&1
ncu -i gpurun_out/${tag}_trace.ncu-rep --page raw --csv > gpurun_out/${tag}_trace_raw.csv
ncu -i gpurun_out/${tag}_trace.ncu-rep --page source --csv > gpurun_out/${tag}_trace_src.csv
python tools/ncu_summary.py gpurun_out/${tag}_trace_raw.csv
