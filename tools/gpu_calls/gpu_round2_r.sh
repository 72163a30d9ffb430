#!/bin/bash
# round-2 GPU call R (2 GPUs): interleaved LiDAR shards — multi-GPU tests, single-GPU LiDAR tests, C4 at N=1 and N=2
mkdir -p gpurun_out
O=gpurun_out
( time timeout 1200 python -m pytest tests/test_gpu_multi.py tests/test_gpu_entrypoints.py tests/test_gpu_configs.py -x -q 2>&1 | tail -15 ) > $O/r2r_pytest.log 2>&1
timeout 600 python bench.py --configs C4 --steps 3 --warmup 3 > $O/r2r_bench_c4_n1.json 2> $O/r2r_bench_c4_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 \
    bench.py --gpus 2 --configs C4 --steps 3 --warmup 3 > $O/r2r_bench_c4_n2.json 2> $O/r2r_bench_c4_n2.err
cat $O/r2r_pytest.log; tail -n 3 $O/r2r_bench_c4_n1.err $O/r2r_bench_c4_n2.err
python - <<PY
import json
for n in (1,2):
    try:
        d=json.load(open(f"gpurun_out/r2r_bench_c4_n{n}.json")); c=d["configs"]["C4"]
        print(n, "C4", c["ms"], c["ms_incl_d2h"], c["Mrays_per_s"], c.get("parity",{}).get("ok"), c["trace_ms_per_rank"], c["what"])
    except Exception as e: print(n, "failed", e)
PY
