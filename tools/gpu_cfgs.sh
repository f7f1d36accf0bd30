#!/bin/bash
# GPU box: the BASELINE configurations at a reduced spp (default form per scene, forced staged, forced fused).
O=gpurun_out; mkdir -p $O
for wl in terrain caustics spheres cornell; do
  spp=32; [ $wl = terrain ] && spp=8
  echo "== $wl default"; python tools/prof.py --workload $wl --spp $spp --sample-count 128 --renders 3 2>&1 | tail -1
  echo "== $wl MTS_SHADE_FUSED=1"; MTS_SHADE_FUSED=1 python tools/prof.py --workload $wl --spp $spp --sample-count 128 --renders 3 2>&1 | tail -1
  echo "== $wl MTS_SHADE_STAGED=1"; MTS_SHADE_STAGED=1 python tools/prof.py --workload $wl --spp $spp --sample-count 128 --renders 3 2>&1 | tail -1
done | tee $O/${1:-cfgs}_cfgs.txt
python -m pytest tests -q -m gpu 2>&1 | tail -2
