#!/bin/bash
# A/B timing of library variants at n=$2: bash tools/gpu_pk_ab.sh TAG N "default fu84 ..."   (variants: lib/ab/libtlfea_NAME.so)
TAG=${1:-ab}; N=${2:-119}; VARS=${3:-"default"}
mkdir -p gpurun_out
for v in $VARS; do
  echo "== variant $v" >> gpurun_out/${TAG}_ab.log
  if [ "$v" = default ]; then unset TLFEA_LIB; else export TLFEA_LIB=$PWD/total-lagrangian-fea_b200/lib/ab/libtlfea_$v.so; fi
  timeout 300 python tools/time_pk.py $N 4 2 >> gpurun_out/${TAG}_ab.log 2>&1
done
grep -E "^==|ms/assemble" gpurun_out/${TAG}_ab.log | cut -c1-150
