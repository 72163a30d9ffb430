#!/bin/bash
# round-2 GPU call A: parity suite on the new node format + traversal variants
mkdir -p gpurun_out
V=raytracerfacility_b200/variants
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/r2a_pytest.log 2>&1
for v in default ffma2 ffma2_80 r1; do
  if [ $v = default ]; then unset RTF_B200_LIB; else export RTF_B200_LIB=$PWD/$V/librtf_$v.so; fi
  chk=131072; [ $v = r1 ] && chk=0
  echo "== $v trace" >> gpurun_out/r2a_variants.log
  timeout 300 python tools/quick_trace.py --reps 5 --check $chk >> gpurun_out/r2a_variants.log 2>&1
  echo "== $v illum" >> gpurun_out/r2a_variants.log
  timeout 300 python tools/quick_illum.py --reps 2 >> gpurun_out/r2a_variants.log 2>&1
done
unset RTF_B200_LIB
timeout 600 bash tools/ncu_trace.sh r2a > gpurun_out/r2a_ncu_trace.log 2>&1
rm -f gpurun_out/r2a_trace.ncu-rep
cat gpurun_out/r2a_pytest.log gpurun_out/r2a_variants.log
