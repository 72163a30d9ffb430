#!/bin/bash
# round-2 GPU call I: wide (level-by-level) BTF evaluator A/B on C2 / C5, parity tests of the material paths
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_materials.py tests/test_gpu_configs.py -x -q 2>&1 | tail -8 ) > gpurun_out/r2i_pytest.log 2>&1
V=raytracerfacility_b200/variants
for v in default btfnarrow btfwide3; do
  if [ $v = default ]; then unset RTF_B200_LIB; else export RTF_B200_LIB=$PWD/$V/librtf_$v.so; fi
  timeout 600 python bench.py --configs C2,C5 --steps 3 --warmup 3 > gpurun_out/r2i_bench_$v.json 2> gpurun_out/r2i_bench_$v.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r2i_bench_$v.json"))
print("$v", "M1", round(d["value"],1), {k:(round(c["ms"],2), {p:round(x,2) for p,x in c["phases_ms_rank0"].items() if x}, c["parity"]["ok"]) for k,c in d["configs"].items()})
PY
done > gpurun_out/r2i_variants.log 2>&1
unset RTF_B200_LIB
cat gpurun_out/r2i_pytest.log gpurun_out/r2i_variants.log
