#!/bin/bash
# node-quorum sweep on the v2 node format (M1 and the 1 M x 64 illumination job)
mkdir -p gpurun_out
O=gpurun_out
V=raytracerfacility_b200/variants
for v in quorum16 quorum20 default quorum28; do
  if [ $v = default ]; then unset RTF_B200_LIB; else export RTF_B200_LIB=$PWD/$V/librtf_$v.so; fi
  echo "== $v" >> $O/r2_quorum.log
  timeout 200 python tools/quick_trace.py --reps 5 --check 0 >> $O/r2_quorum.log 2>&1
  timeout 200 python tools/quick_illum.py --reps 2 >> $O/r2_quorum.log 2>&1
done
cat $O/r2_quorum.log
