#!/bin/bash
# round-2 GPU call N (2 GPUs): in-library multi-GPU tests (C host, one process; torchrun ranks) and the bench at N=2
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi -L > $O/r2n_gpus.log 2>&1
( time timeout 900 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -8 ) > $O/r2n_pytest.log 2>&1
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 \
    bench.py --gpus 2 --steps 5 --warmup 3 > $O/r2n_bench_n2.json 2> $O/r2n_bench_n2.err ) 2> $O/r2n_bench_n2.time
cat $O/r2n_gpus.log $O/r2n_pytest.log $O/r2n_bench_n2.time; tail -n 4 $O/r2n_bench_n2.err; head -c 600 $O/r2n_bench_n2.json
