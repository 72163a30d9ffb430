#!/bin/bash
# round-2 GPU call V: sort tile-size / occupancy variants (build timings, sort parity), LCG test
mkdir -p gpurun_out
O=gpurun_out
V=raytracerfacility_b200/variants
for v in default sort12_4 sort8_5 sort8_6; do
  if [ $v = default ]; then unset RTF_B200_LIB; else export RTF_B200_LIB=$PWD/$V/librtf_$v.so; fi
  echo "== $v" >> $O/r2v_sort.log
  timeout 300 python tools/quick_build.py --reps 3 2>&1 | grep -E "rebuild" | tail -1 >> $O/r2v_sort.log
  timeout 300 python tools/quick_build.py --plants 500 --reps 2 2>&1 | grep -E "rebuild" | tail -1 >> $O/r2v_sort.log
  timeout 300 python -m pytest tests/test_gpu_build.py -x -q 2>&1 | tail -1 >> $O/r2v_sort.log
done
unset RTF_B200_LIB
timeout 600 python -m pytest tests/test_gpu_lcg.py tests/test_gpu_entrypoints.py -x -q 2>&1 | tail -3 >> $O/r2v_sort.log
cat $O/r2v_sort.log
