#!/bin/bash
# time the patch assembly at n=$2 under a list of TLFEA_PK_DEBUG masks: bash tools/gpu_pk_dbg.sh TAG N "0 16 32 64 96"
TAG=${1:-r2q}; N=${2:-119}; MASKS=${3:-"0 16 32 64 96"}
mkdir -p gpurun_out
for dbg in $MASKS; do
  echo "== TLFEA_PK_DEBUG=$dbg" >> gpurun_out/${TAG}_phases.log
  TLFEA_PK_DEBUG=$dbg timeout 300 python tools/time_pk.py $N 4 2 >> gpurun_out/${TAG}_phases.log 2>&1
done
grep -E "^==|ms/assemble" gpurun_out/${TAG}_phases.log | cut -c1-150
