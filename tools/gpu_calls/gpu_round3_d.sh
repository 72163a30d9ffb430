#!/bin/bash
# round 3, call d: determinism of the hand-synchronised kernels (the sanitizer is closed on this pool), the full bench line
# with per-leg rooflines, and ncu --set full of k_trace_queue<0> on the C3 scene (10 M triangles), bounces 0..4 of one batch
mkdir -p gpurun_out
python tools/sanitize_workload.py > gpurun_out/r3d_workload_plain.log 2>&1; echo "workload rc=$?"
python tools/determinism_check.py --plants 50 --reps 12 > gpurun_out/r3d_determinism_c2.log 2>&1; echo "det50 rc=$?"; tail -n 2 gpurun_out/r3d_determinism_c2.log
python tools/determinism_check.py --plants 500 --reps 4 > gpurun_out/r3d_determinism_c3.log 2>&1; echo "det500 rc=$?"; tail -n 2 gpurun_out/r3d_determinism_c3.log
python bench.py > gpurun_out/r3d_bench.json 2> gpurun_out/r3d_bench.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/r3d_bench.json
Q="python tools/quick_illum.py --plants 500 --samples 32 --tris 131072 --tri-begin 5000000 --reps 1"
$Q > gpurun_out/r3d_c3_plain.log 2>&1; echo "quick rc=$?"; cat gpurun_out/r3d_c3_plain.log | tail -n 2
ncu --set full --clock-control none --import-source on -k regex:k_trace_queue -s 0 -c 5 -f -o gpurun_out/r3d_c3_trace $Q > gpurun_out/r3d_c3_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r3d_c3_trace.ncu-rep --page raw --csv > gpurun_out/r3d_c3_trace_raw.csv
ncu -i gpurun_out/r3d_c3_trace.ncu-rep --page source --csv > gpurun_out/r3d_c3_trace_src.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/r3d_c3_trace_raw.csv > gpurun_out/r3d_c3_trace_summary.txt 2>&1; head -n 60 gpurun_out/r3d_c3_trace_summary.txt
ls -la gpurun_out/ | head -30
