#!/bin/bash
# round-2 GPU call K: register-only collapse plan + batched sort look-back (parity + timings), BSSRDF shade occupancy variants
mkdir -p gpurun_out
O=gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_build.py tests/test_gpu_trace.py tests/test_gpu_prebake.py -x -q 2>&1 | tail -8 ) > $O/r2k_pytest.log 2>&1
timeout 300 python tools/quick_build.py --reps 3 > $O/r2k_build.log 2>&1
timeout 300 python tools/quick_build.py --plants 500 --reps 3 > $O/r2k_build_c3.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/r2k_build_launches.csv python tools/quick_build.py --reps 1 > /dev/null 2>&1
V=raytracerfacility_b200/variants
for v in default sss2 sss3; do
  if [ $v = default ]; then unset RTF_B200_LIB; else export RTF_B200_LIB=$PWD/$V/librtf_$v.so; fi
  echo "== $v" >> $O/r2k_sss.log
  timeout 300 python tools/quick_camera.py --config C2 --spp 16 --sss 1 >> $O/r2k_sss.log 2>&1
done
unset RTF_B200_LIB
timeout 600 python -m pytest tests/test_gpu_sss.py -x -q 2>&1 | tail -3 >> $O/r2k_sss.log
cat $O/r2k_pytest.log $O/r2k_build.log $O/r2k_build_c3.log $O/r2k_sss.log
