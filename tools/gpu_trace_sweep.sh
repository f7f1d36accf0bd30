#!/bin/bash
# GPU box: traversal tuning sweep (build/libmtscuda_*.so variants of the trace kernels) + the L2 persisting-window A/B.
TAG=${1:-sweep}; O=gpurun_out; mkdir -p $O
{
echo "== base"; python tools/prof.py --spp 8 --sample-count 128 --renders 3 2>&1 | tail -2
echo "== MTS_L2_WINDOW=1"; MTS_L2_WINDOW=1 python tools/prof.py --spp 8 --sample-count 128 --renders 3 2>&1 | tail -2
for f in build/libmtscuda_*.so; do
  [ -e "$f" ] || continue
  echo "== $f"; MTS_CUDA_LIB=$PWD/$f python tools/prof.py --spp 8 --sample-count 128 --renders 3 2>&1 | tail -2
done
echo "== base spp 128"; python tools/prof.py --spp 128 --renders 2 2>&1 | tail -1
echo "== MTS_L2_WINDOW=1 spp 128"; MTS_L2_WINDOW=1 python tools/prof.py --spp 128 --renders 2 2>&1 | tail -1
} | tee $O/${TAG}_sweep.txt
for w in 0 1; do
  if [ $w = 1 ]; then export MTS_L2_WINDOW=1; fi
  ncu --metrics gpu__time_duration.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,dram__bytes_read.sum --clock-control none -k regex:"k_intersect|k_shadow" -c 6 --csv --log-file $O/${TAG}_l2window${w}.csv python tools/prof.py --spp 8 --sample-count 128 > /dev/null 2>&1
done
