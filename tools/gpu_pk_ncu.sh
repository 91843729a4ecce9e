#!/bin/bash
# ONE ncu --set full capture (with source) of two mid-sequence pk_kernel launches at n=64 and its csv pages.
TAG=${1:-r2w}
mkdir -p gpurun_out
timeout 900 ncu --set full --import-source on --clock-control none -k regex:pk_kernel -s 12 -c 2 \
  -o gpurun_out/${TAG}_pk_full -f python tools/time_pk.py 64 4 2 > gpurun_out/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i gpurun_out/${TAG}_pk_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_pk_raw.csv 2>/dev/null
ncu -i gpurun_out/${TAG}_pk_full.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/${TAG}_pk_source.csv 2>/dev/null
ncu -i gpurun_out/${TAG}_pk_full.ncu-rep --page details > gpurun_out/${TAG}_pk_details.txt 2>/dev/null
python tools/ncu_metrics.py gpurun_out/${TAG}_pk_raw.csv > gpurun_out/${TAG}_pk_metrics.txt 2>&1
python tools/ncu_regions.py gpurun_out/${TAG}_pk_source.csv > gpurun_out/${TAG}_pk_regions.txt 2>&1
cat gpurun_out/${TAG}_pk_metrics.txt gpurun_out/${TAG}_pk_regions.txt
