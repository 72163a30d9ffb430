#!/bin/bash
# round-2 GPU call X (final): full GPU suite, full bench of both arms, launch list, ncu capture of k_trace_rays (traffic)
mkdir -p gpurun_out
O=gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 ) > $O/r2x_pytest.log 2>&1
( time timeout 1500 python bench.py > $O/r2x_bench.json 2> $O/r2x_bench.err ) 2> $O/r2x_bench.time
( time timeout 900 python bench.py --impl reference > $O/r2x_bench_ref.json 2> $O/r2x_bench_ref.err ) 2> $O/r2x_bench_ref.time
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r2x_bench_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-extras > $O/r2x_bench_ncu.log 2>&1
timeout 900 bash tools/ncu_trace.sh r2x > $O/r2x_ncu_trace.log 2>&1
rm -f $O/r2x_trace.ncu-rep
cat $O/r2x_pytest.log $O/r2x_bench.time $O/r2x_bench_ref.time; tail -n 3 $O/r2x_bench.err; tail -n 3 $O/r2x_bench_ref.err; tail -n 12 $O/r2x_ncu_trace.log
