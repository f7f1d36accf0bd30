#!/bin/bash
# One GPU-box call: parity tests, smoke, a short bench run, then (only after those exited 0 without a profiler) the ncu
# launch list and one `ncu --set full` capture of the hot kernels.  Usage: tools/gpu_batch.sh <tag> [bench args]
TAG=${1:-r2}; shift
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest.log
tail -5 $O/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/${TAG}_smoke.log
python bench.py --steps 3 --warmup 3 "$@" > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; tail -c 600 $O/${TAG}_bench.err
MTS_TRACE_BOUNCES=1 python tools/prof.py --spp 8 --sample-count 128 > $O/${TAG}_plain.log 2> $O/${TAG}_bounce_trace.txt; echo "prof rc=$?"; cat $O/${TAG}_plain.log
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_raygen|k_intersect|k_shadow|k_shade|k_sort|k_splat|k_ewa|k_begin" -c 400 --csv --log-file $O/${TAG}_launches.csv python tools/prof.py --spp 8 --sample-count 128 > $O/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_intersect|k_shade|k_shadow" -c 32 -f -o /tmp/${TAG}_full python tools/prof.py --spp 8 --sample-count 128 > $O/${TAG}_ncu2.log 2>&1
ncu -i /tmp/${TAG}_full.ncu-rep --page raw --csv > $O/${TAG}_full_raw.csv 2>/dev/null
ncu -i /tmp/${TAG}_full.ncu-rep --page source --csv --print-source cuda,sass -k regex:k_shade -c 2 > /tmp/${TAG}_src.csv 2>/dev/null
python tools/ncu_lines.py /tmp/${TAG}_src.csv > $O/${TAG}_shade_lines.txt 2>/dev/null
ls -la $O | tail -12
