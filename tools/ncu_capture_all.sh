#!/bin/bash
# `ncu --set full` over every launch of one (short) render per workload; raw CSV pages + the segment counts of the same render
# into gpurun_out/ncu_$1/.  usage (GPU box): bash tools/ncu_capture_all.sh <tag> [scene:spp ...]
TAG=${1:-r02z}; shift
SPECS=${@:-pawn:32 bunny_volume:24 demo1:24 demo2:24 cornell:200 cornell_noredirect:200 readme:50}
OUT=gpurun_out/ncu_$TAG; mkdir -p $OUT
for spec in $SPECS; do
  name=${spec%%:*}; spp=${spec##*:}
  python tools/gpu_traffic_pass.py $name $spp stats > $OUT/${name}_stats.json 2> $OUT/${name}_stats.err || { echo "stats pass failed: $name"; continue; }
  timeout 600 ncu --set full --clock-control none -k regex:'k_traverse_p|k_shade|k_raygen|k_trace_flat|k_finish' -f -o /tmp/ncu_$name \
      python tools/gpu_traffic_pass.py $name $spp plain > $OUT/${name}_ncu.log 2>&1 || { echo "ncu pass failed: $name"; tail -3 $OUT/${name}_ncu.log; continue; }
  ncu -i /tmp/ncu_$name.ncu-rep --page raw --csv > $OUT/${name}_raw.csv 2>> $OUT/${name}_ncu.log
  echo "$name: $(cat $OUT/${name}_stats.json) rows $(wc -l < $OUT/${name}_raw.csv)"
done
