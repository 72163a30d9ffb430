#!/bin/bash
# Launch list of the bench command (after it exited 0 without ncu) + full capture of the dominant kernel.
# usage: tools/profile_bench.sh <tag>
set -e
tag=${1:-cur}
python bench.py --steps 2 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_bench_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/${tag}_bench_ncu.log 2>&1
bash tools/ncu_trace.sh ${tag}
