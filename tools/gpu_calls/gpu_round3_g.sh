#!/bin/bash
# round 3, call g (N GPUs): headline + e2e + copy-only host ceiling only (--no-extras), to see where e2e saturates
N=${1:-4}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --no-extras \
    > gpurun_out/r3g_bench_n$N.json 2> gpurun_out/r3g_bench_n$N.err; echo "bench N=$N rc=$?"
python - <<PY
import json
b = json.load(open("gpurun_out/r3g_bench_n$N.json"))
print("value", b["value"], "e2e", json.dumps(b["e2e"]))
PY
nvidia-smi topo -m > gpurun_out/r3g_topo_n$N.txt 2>&1; lscpu | grep -i "numa\|socket\|model name" > gpurun_out/r3g_cpu_n$N.txt
