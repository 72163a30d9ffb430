#!/bin/bash
# round-2 GPU call J: ncu of k_shade (BTF wavefront of C5; BSSRDF variant), ncu of the builder kernels, trace variants
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python tools/quick_camera.py --config C5 --spp 8 > $O/r2j_c5_plain.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 6 -c 1 -f -o $O/r2j_shade_btf \
    python tools/quick_camera.py --config C5 --spp 8 --reps 1 > $O/r2j_shade_btf_ncu.log 2>&1
ncu -i $O/r2j_shade_btf.ncu-rep --page raw --csv > $O/r2j_shade_btf_raw.csv
ncu -i $O/r2j_shade_btf.ncu-rep --page source --csv > $O/r2j_shade_btf_src.csv
rm -f $O/r2j_shade_btf.ncu-rep
timeout 300 python tools/quick_camera.py --config C2 --spp 16 --sss 1 > $O/r2j_sss_plain.log 2>&1
timeout 300 python tools/quick_camera.py --config C2 --spp 16 --sss 0 >> $O/r2j_sss_plain.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 6 -c 1 -f -o $O/r2j_shade_sss \
    python tools/quick_camera.py --config C2 --spp 16 --sss 1 --reps 1 > $O/r2j_shade_sss_ncu.log 2>&1
ncu -i $O/r2j_shade_sss.ncu-rep --page raw --csv > $O/r2j_shade_sss_raw.csv
ncu -i $O/r2j_shade_sss.ncu-rep --page source --csv > $O/r2j_shade_sss_src.csv
rm -f $O/r2j_shade_sss.ncu-rep
timeout 300 python tools/quick_build.py --reps 1 > $O/r2j_build_plain.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:'k_collapse|k_sort_pass|k_karras|k_split_emit|k_node_boxes|k_seg_reduce|k_tri_records|k_sort_hist|k_morton|k_flatten|k_bounds' \
    -s 26 -c 26 -f -o $O/r2j_build python tools/quick_build.py --reps 1 > $O/r2j_build_ncu.log 2>&1
ncu -i $O/r2j_build.ncu-rep --page raw --csv > $O/r2j_build_raw.csv
rm -f $O/r2j_build.ncu-rep
V=raytracerfacility_b200/variants
for v in ffma2 queue8; do
  export RTF_B200_LIB=$PWD/$V/librtf_$v.so
  echo "== $v trace" >> $O/r2j_variants.log
  timeout 300 python tools/quick_trace.py --reps 5 --check 131072 >> $O/r2j_variants.log 2>&1
  echo "== $v illum" >> $O/r2j_variants.log
  timeout 300 python tools/quick_illum.py --reps 2 >> $O/r2j_variants.log 2>&1
done
unset RTF_B200_LIB
cat $O/r2j_c5_plain.log $O/r2j_sss_plain.log $O/r2j_build_plain.log $O/r2j_variants.log
tail -3 $O/r2j_shade_btf_ncu.log $O/r2j_shade_sss_ncu.log $O/r2j_build_ncu.log
