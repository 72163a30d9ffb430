#!/bin/bash
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 ) > gpurun_out/r2b_pytest.log 2>&1
timeout 300 python tools/quick_trace.py --reps 5 > gpurun_out/r2b_trace.log 2>&1
timeout 300 python tools/quick_illum.py --reps 2 >> gpurun_out/r2b_trace.log 2>&1
timeout 600 python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err
timeout 600 bash tools/ncu_trace.sh r2b > gpurun_out/r2b_ncu_trace.log 2>&1
rm -f gpurun_out/r2b_trace.ncu-rep
cat gpurun_out/r2b_pytest.log gpurun_out/r2b_trace.log; tail -3 gpurun_out/r2b_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2b_bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','build','rebuild','refit')})
PY
