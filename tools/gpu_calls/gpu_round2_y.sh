#!/bin/bash
# round-2 GPU call Y (8 GPUs): multi-GPU tests and C4 / C1 with the finer LiDAR interleave
mkdir -p gpurun_out
O=gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -5 ) > $O/r2y_pytest.log 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 \
    bench.py --gpus 8 --configs C4 --steps 5 --warmup 3 > $O/r2y_bench_c4_n8.json 2> $O/r2y_bench_c4_n8.err
cat $O/r2y_pytest.log; tail -n 2 $O/r2y_bench_c4_n8.err
python - <<PY
import json
d=json.load(open("gpurun_out/r2y_bench_c4_n8.json")); c=d["configs"]["C4"]
print("C4 n8", c["ms"], c["ms_incl_d2h"], c["Mrays_per_s"], c["trace_ms_per_rank"], c["what"])
PY
