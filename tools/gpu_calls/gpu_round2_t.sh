#!/bin/bash
# round-2 GPU call T (2 GPUs): camera bands of a rank as one interleaved job — parity tests, C2/C5 at N=1 and N=2
mkdir -p gpurun_out
O=gpurun_out
( time timeout 1500 python -m pytest tests/test_gpu_multi.py tests/test_gpu_entrypoints.py tests/test_gpu_configs.py -x -q 2>&1 | tail -15 ) > $O/r2t_pytest.log 2>&1
timeout 600 python bench.py --configs C2,C5 --steps 3 --warmup 3 > $O/r2t_bench_n1.json 2> $O/r2t_bench_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 \
    bench.py --gpus 2 --configs C2,C5 --steps 3 --warmup 3 > $O/r2t_bench_n2.json 2> $O/r2t_bench_n2.err
cat $O/r2t_pytest.log; tail -n 3 $O/r2t_bench_n1.err $O/r2t_bench_n2.err
python - <<PY
import json
for n in (1,2):
    try:
        d=json.load(open(f"gpurun_out/r2t_bench_n{n}.json"))
        for k,c in d["configs"].items():
            print(n, k, round(c["ms"],2), {a:round(b,2) for a,b in c["phases_ms_rank0"].items() if b}, c.get("parity",{}).get("ok"), c.get("image_mean"))
    except Exception as e: print(n, "failed", e)
PY
