#!/bin/bash
# GPU box: parity tests with the regular build, then tools/prof.py timings of env / build variants, then ncu of the shade kernels.
# Usage: tools/gpu_variants.sh <tag> [spp]
TAG=${1:-var}; SPP=${2:-8}
O=gpurun_out; mkdir -p $O
python -m pytest tests -q -m gpu > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
{
echo "== base"; python tools/prof.py --spp $SPP --sample-count 128 --renders 3 2>&1 | tail -2
echo "== MTS_EWA_INLINE=1"; MTS_EWA_INLINE=1 python tools/prof.py --spp $SPP --sample-count 128 --renders 3 2>&1 | tail -2
echo "== MTS_SHADE_FUSED=1"; MTS_SHADE_FUSED=1 python tools/prof.py --spp $SPP --sample-count 128 --renders 3 2>&1 | tail -2
for f in build/libmtscuda_*.so; do
  [ -e "$f" ] || continue
  echo "== $f"; MTS_CUDA_LIB=$PWD/$f python tools/prof.py --spp $SPP --sample-count 128 --renders 3 2>&1 | tail -2
done
echo "== base spp 32"; python tools/prof.py --spp 32 --renders 2 2>&1 | tail -1
echo "== base spp 128"; python tools/prof.py --spp 128 --renders 2 2>&1 | tail -1
} | tee $O/${TAG}_variants.txt
MTS_TRACE_BOUNCES=1 python tools/prof.py --spp 8 --sample-count 128 > /dev/null 2> $O/${TAG}_bounce_trace.txt
ncu --set full --clock-control none --import-source on -k regex:"k_shade_surface|k_shade_scatter|k_shade_ewa" -c 5 -f -o /tmp/${TAG}_full python tools/prof.py --spp 8 --sample-count 128 > $O/${TAG}_ncu2.log 2>&1
ncu -i /tmp/${TAG}_full.ncu-rep --page raw --csv > $O/${TAG}_full_raw.csv 2>/dev/null
ncu -i /tmp/${TAG}_full.ncu-rep --page source --csv --print-source cuda,sass -k regex:k_shade_surface -c 1 > /tmp/${TAG}_src1.csv 2>/dev/null
python tools/ncu_lines.py /tmp/${TAG}_src1.csv > $O/${TAG}_surface_lines.txt 2>/dev/null
ncu -i /tmp/${TAG}_full.ncu-rep --page source --csv --print-source cuda,sass -k regex:k_shade_ewa -c 1 > /tmp/${TAG}_src2.csv 2>/dev/null
python tools/ncu_lines.py /tmp/${TAG}_src2.csv > $O/${TAG}_ewa_lines.txt 2>/dev/null
