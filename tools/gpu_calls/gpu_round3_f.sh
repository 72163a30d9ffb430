#!/bin/bash
# round 3, call f (N GPUs): the bench line under torchrun as the driver launches it (host ceiling of e2e, multi_gpu_parity, per-leg rooflines)
N=${1:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N \
    > gpurun_out/r3f_bench_n$N.json 2> gpurun_out/r3f_bench_n$N.err; echo "bench N=$N rc=$?"
tail -c 1800 gpurun_out/r3f_bench_n$N.json; tail -n 5 gpurun_out/r3f_bench_n$N.err
