#!/bin/bash
# bench at N GPUs through torchrun (the driver's launch line) + the multi-GPU tests; usage: gpu_round2_scale.sh N
N=${1:-8}
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi -L > $O/r2s_gpus_n$N.log 2>&1
( time timeout 900 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -5 ) > $O/r2s_pytest_n$N.log 2>&1
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 \
    bench.py --gpus $N --steps 5 --warmup 3 > $O/r2s_bench_n$N.json 2> $O/r2s_bench_n$N.err ) 2> $O/r2s_bench_n$N.time
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 \
    bench.py --impl reference --gpus $N --steps 3 --warmup 1 > $O/r2s_bench_ref_n$N.json 2> $O/r2s_bench_ref_n$N.err ) 2> $O/r2s_bench_ref_n$N.time
cat $O/r2s_pytest_n$N.log $O/r2s_bench_n$N.time; tail -n 3 $O/r2s_bench_n$N.err; wc -l $O/r2s_bench_n$N.json $O/r2s_bench_ref_n$N.json; head -c 300 $O/r2s_bench_n$N.json
