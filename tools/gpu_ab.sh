#!/bin/bash
# A/B of the staged (k_surface + k_scatter) and fused (k_shade<leaf>) forms on the GPU box + ncu of the staged kernels.
# Usage: tools/gpu_ab.sh <tag>
TAG=${1:-ab}
O=gpurun_out
mkdir -p $O
python -m pytest tests -q -m gpu > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/${TAG}_pytest.log
for spp in 8 32; do
  echo "== staged spp $spp"; python tools/prof.py --spp $spp --renders 3 2>&1 | tail -2
  echo "== fused spp $spp"; MTS_SHADE_FUSED=1 python tools/prof.py --spp $spp --renders 3 2>&1 | tail -2
done | tee $O/${TAG}_ab.txt
MTS_TRACE_BOUNCES=1 python tools/prof.py --spp 8 > /dev/null 2> $O/${TAG}_bounce_trace.txt
ncu --set full --clock-control none --import-source on -k regex:"k_shade_surface|k_shade_scatter|k_shade<" -c 6 -f -o /tmp/${TAG}_full python tools/prof.py --spp 8 > $O/${TAG}_ncu2.log 2>&1
ncu -i /tmp/${TAG}_full.ncu-rep --page raw --csv > $O/${TAG}_full_raw.csv 2>/dev/null
ncu -i /tmp/${TAG}_full.ncu-rep --page source --csv --print-source cuda,sass -k regex:k_shade_surface -c 1 > /tmp/${TAG}_src1.csv 2>/dev/null
python tools/ncu_lines.py /tmp/${TAG}_src1.csv > $O/${TAG}_surface_lines.txt 2>/dev/null
ncu -i /tmp/${TAG}_full.ncu-rep --page source --csv --print-source cuda,sass -k regex:k_shade_scatter -c 1 > /tmp/${TAG}_src2.csv 2>/dev/null
python tools/ncu_lines.py /tmp/${TAG}_src2.csv > $O/${TAG}_scatter_lines.txt 2>/dev/null
ls -la $O | tail -8
