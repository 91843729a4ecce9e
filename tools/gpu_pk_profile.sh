#!/bin/bash
# pk_kernel deep dive on one B200: phase breakdown by the TLFEA_PK_DEBUG switches at n=119, then ONE ncu --set full
# capture (with source) of a mid-sequence pk_kernel launch at n=64 and its csv pages.  Outputs under gpurun_out/.
TAG=${1:-r2p}
N=${2:-119}
mkdir -p gpurun_out
for dbg in 0 1 3 7 15 8 4 2; do
  echo "== TLFEA_PK_DEBUG=$dbg" >> gpurun_out/${TAG}_phases.log
  TLFEA_PK_DEBUG=$dbg timeout 300 python tools/time_pk.py $N 4 2 >> gpurun_out/${TAG}_phases.log 2>&1
done
grep -E "^==|ms/assemble" gpurun_out/${TAG}_phases.log | cut -c1-220
# ncu: skip the first colour launches of the warm-up assembles, capture 2 launches
timeout 900 ncu --set full --import-source on --clock-control none -k regex:pk_kernel -s 12 -c 2 \
  -o gpurun_out/${TAG}_pk_full -f python tools/time_pk.py 64 4 2 > gpurun_out/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i gpurun_out/${TAG}_pk_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_pk_raw.csv 2>/dev/null
ncu -i gpurun_out/${TAG}_pk_full.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/${TAG}_pk_source.csv 2>/dev/null
ncu -i gpurun_out/${TAG}_pk_full.ncu-rep --page details > gpurun_out/${TAG}_pk_details.txt 2>/dev/null
python tools/ncu_metrics.py gpurun_out/${TAG}_pk_raw.csv > gpurun_out/${TAG}_pk_metrics.txt 2>&1
python tools/ncu_regions.py gpurun_out/${TAG}_pk_source.csv > gpurun_out/${TAG}_pk_regions.txt 2>&1
cat gpurun_out/${TAG}_pk_regions.txt
ls -la gpurun_out | head -30
