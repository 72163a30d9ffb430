#!/bin/bash
# round-2 GPU call H: parity suite after the stats fix, pipe micro-benchmark, refill-threshold variants
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/r2h_pytest.log 2>&1
timeout 120 tools/ubench/pipes > gpurun_out/r2h_pipes.log 2>&1
V=raytracerfacility_b200/variants
for v in default refill2 refill3 refill4; do
  if [ $v = default ]; then unset RTF_B200_LIB; else export RTF_B200_LIB=$PWD/$V/librtf_$v.so; fi
  echo "== $v trace" >> gpurun_out/r2h_variants.log
  timeout 300 python tools/quick_trace.py --reps 5 --check 131072 >> gpurun_out/r2h_variants.log 2>&1
  echo "== $v illum" >> gpurun_out/r2h_variants.log
  timeout 300 python tools/quick_illum.py --reps 2 >> gpurun_out/r2h_variants.log 2>&1
done
unset RTF_B200_LIB
cat gpurun_out/r2h_pytest.log gpurun_out/r2h_pipes.log gpurun_out/r2h_variants.log
