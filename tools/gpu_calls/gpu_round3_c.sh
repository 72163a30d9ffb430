#!/bin/bash
# round 3, call c: compute-sanitizer (memcheck, synccheck, initcheck, racecheck) over tools/sanitize_workload.py
bash tools/sanitize.sh r3c
cat gpurun_out/r3c_summary.txt
