#!/bin/bash
# Regenerates the evidence under profiles/ for one round: run under gpurun, ONE ncu mode per call
# (B200_PROFILING.md).  usage: tools/profile_round.sh <tag> <what>
#   what = launches   per-launch device times of the C2 / C3 / wide (C2-ROT shape) paths
#        = full-c2    ncu --set full of k_batch
#        = full-c3    ncu --set full of k_sweep, k_f2, k_f1
#        = full-wide  ncu --set full of the four wide kernels (first chunk round of frame 2)
# Each plain command runs first (no ncu); summaries: tools/ncu_summary.py, tools/ncu_lines.py, tools/launch_summary.py.
set -e
TAG=${1:-rXX}; WHAT=${2:-launches}
mkdir -p gpurun_out
C2="python bench.py --seqs 444 --frames 100 --steps 1 --warmup 1 --no-extras --cpu-seconds 1"
C3="python bench.py --only c3"
WIDE="python tools/wide_bench.py rot 4096 32 6 0.5"
case $WHAT in
launches)
  $C2 > gpurun_out/plain_c2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/${TAG}_c2_launches.csv $C2 > /dev/null 2>&1
  $C3 > gpurun_out/plain_c3.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 400 -c 45 --csv --log-file gpurun_out/${TAG}_c3_launches.csv $C3 > /dev/null 2>&1
  $WIDE > gpurun_out/plain_wide.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_wide_launches.csv $WIDE > /dev/null 2>&1 ;;
full-c2)
  $C2 > gpurun_out/plain_c2.log 2>&1 && ncu --set full --clock-control none --import-source on -k "regex:^k_batch$" -c 1 -f -o gpurun_out/${TAG}_k_batch $C2 > /dev/null 2>&1 ;;
full-c3)
  $C3 > gpurun_out/plain_c3.log 2>&1 && for k in k_sweep k_f2 k_f1; do ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 150 -c 1 -f -o gpurun_out/${TAG}_$k $C3 > /dev/null 2>&1; done ;;
full-wide)
  $WIDE > gpurun_out/plain_wide.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_wide -s 16 -c 4 -f -o gpurun_out/${TAG}_wide $WIDE > /dev/null 2>&1 ;;
esac
ls -la gpurun_out | grep ${TAG} || true
