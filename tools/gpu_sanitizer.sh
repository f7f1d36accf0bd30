#!/bin/bash
# GPU box: compute-sanitizer memcheck + racecheck + initcheck of one small render per shade form.
O=gpurun_out; mkdir -p $O
for tool in memcheck racecheck; do
  for wl in cornell terrain_small; do
    timeout 600 compute-sanitizer --tool $tool --print-limit 20 python tools/prof.py --workload $wl --spp 1 > $O/sanitizer_${tool}_${wl}.log 2>&1
    echo "$tool $wl rc=$?: $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $O/sanitizer_${tool}_${wl}.log | tail -1)"
  done
done
