#!/bin/bash
# compute-sanitizer over tools/sanitize_workload.py, one tool per run; logs under gpurun_out/ (copy the summaries to
# profiles/).  usage: tools/sanitize.sh [tag] [tools...]      default tools: memcheck synccheck initcheck racecheck
tag=${1:-san}; shift
tools=${@:-memcheck synccheck initcheck racecheck}
mkdir -p gpurun_out
compute-sanitizer --version | head -3 > gpurun_out/${tag}_version.txt
python tools/sanitize_workload.py --small > gpurun_out/${tag}_plain.log 2>&1 || { echo "workload fails WITHOUT the sanitizer"; tail -n 20 gpurun_out/${tag}_plain.log; exit 1; }
for t in $tools; do
  extra=""
  [ "$t" = memcheck ] && extra="--leak-check no"
  [ "$t" = initcheck ] && extra="--track-unused-memory no"
  small=""
  [ "$t" = racecheck ] && small="--small"
  /usr/bin/time -f "$t wall %e s" timeout 900 compute-sanitizer --tool $t $extra --print-limit 40 --error-exitcode 9 \
      --log-file gpurun_out/${tag}_$t.log python tools/sanitize_workload.py $small > gpurun_out/${tag}_$t.out 2>&1
  echo "$t rc=$?" | tee -a gpurun_out/${tag}_summary.txt
  tail -n 3 gpurun_out/${tag}_$t.log | tee -a gpurun_out/${tag}_summary.txt
done
