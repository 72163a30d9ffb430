#!/bin/bash
# round-2 GPU call W: block-compacted BTF evaluation — parity tests and C5 / C2 timing
mkdir -p gpurun_out
O=gpurun_out
( time timeout 1500 python -m pytest tests/test_gpu_materials.py tests/test_gpu_configs.py tests/test_gpu_entrypoints.py tests/test_gpu_lcg.py tests/test_gpu_sss.py -x -q 2>&1 | tail -8 ) > $O/r2w_pytest.log 2>&1
timeout 600 python bench.py --configs C2,C5 --steps 3 --warmup 3 > $O/r2w_bench.json 2> $O/r2w_bench.err
cat $O/r2w_pytest.log; tail -n 3 $O/r2w_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/r2w_bench.json"))
for k,c in d["configs"].items():
    print(k, round(c["ms"],2), {a:round(b,2) for a,b in c["phases_ms_rank0"].items() if b}, c.get("parity",{}).get("ok"), c.get("image_mean"))
PY
