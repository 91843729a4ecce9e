#!/bin/bash
# ONE ncu --set full capture (with source) of the ANCF element kernel: bash tools/gpu_aek_ncu.sh TAG 3243 1000000
TAG=${1:-aek}; KIND=${2:-3243}; NX=${3:-1000000}
mkdir -p gpurun_out
timeout 900 ncu --set full --import-source on --clock-control none -k regex:aek_kernel -s 2 -c 1 \
  -o gpurun_out/${TAG}_aek_full -f python tools/time_ancf.py $KIND $NX > gpurun_out/${TAG}_ncu_full.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/${TAG}_aek_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_aek_raw.csv 2>/dev/null
ncu -i gpurun_out/${TAG}_aek_full.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/${TAG}_aek_source.csv 2>/dev/null
python tools/ncu_metrics.py gpurun_out/${TAG}_aek_raw.csv
python tools/ncu_lines.py gpurun_out/${TAG}_aek_source.csv 30 | cut -c1-200
