#!/bin/bash
# 4 GPUs: multi-GPU tests and the sharded LiDAR / camera configurations (band and block counts that do not divide evenly)
mkdir -p gpurun_out
O=gpurun_out
( timeout 600 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -3 ) > $O/r2_n4_pytest.log 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29581 \
    bench.py --gpus 4 --configs C4,C5 --steps 5 --warmup 3 > $O/r2_n4_bench.json 2> $O/r2_n4_bench.err
cat $O/r2_n4_pytest.log; tail -n 2 $O/r2_n4_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/r2_n4_bench.json"))
print("M1", d["value"], "e2e", d["e2e"]["value"])
for k,c in d["configs"].items(): print(k, c["ms"], c["parity"]["ok"], c.get("trace_ms_per_rank"))
PY
