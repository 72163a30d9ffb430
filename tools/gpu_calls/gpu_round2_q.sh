#!/bin/bash
# round-2 GPU call Q: SASS-level profile of k_collapse; C4 after the scan kernel's launch bounds
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python bench.py --configs C4 --steps 3 --warmup 3 > $O/r2q_bench_c4.json 2> $O/r2q_bench_c4.err
python - <<PY
import json
d=json.load(open("gpurun_out/r2q_bench_c4.json"))
c=d["configs"]["C4"]; print("C4", c["ms"], c["Mrays_per_s"], c["parity"]["ok"], "M1", d["value"])
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_collapse' -s 1 -c 1 -f -o $O/r2q_collapse \
    python tools/quick_build.py --reps 1 > $O/r2q_collapse_ncu.log 2>&1
ncu -i $O/r2q_collapse.ncu-rep --page source --csv > $O/r2q_collapse_src.csv
ncu -i $O/r2q_collapse.ncu-rep --page raw --csv > $O/r2q_collapse_raw.csv
rm -f $O/r2q_collapse.ncu-rep
tail -n 2 $O/r2q_collapse_ncu.log
