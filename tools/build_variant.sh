#!/bin/bash
# A/B build of libtlfea_b200.so with extra nvcc flags:  tools/build_variant.sh NAME "-DPK_FU_OWN=8"
# -> total-lagrangian-fea_b200/lib/ab/libtlfea_NAME.so  (select it at run time with TLFEA_LIB=...)
set -e
cd "$(dirname "$0")/../total-lagrangian-fea_b200"
name=$1; extra=$2
bd=build_ab/$name; mkdir -p $bd lib/ab
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fopenmp -ccbin /usr/bin/g++ --expt-relaxed-constexpr"
pids=()
for f in elem newton pattern ancf order patch_plan patch_asm dist; do $NV $extra -c csrc/$f.cu -o $bd/$f.o & pids+=($!); done
for p in "${pids[@]}"; do wait $p; done
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o lib/ab/libtlfea_$name.so $bd/*.o -lcudart -lgomp -ldl
echo built lib/ab/libtlfea_$name.so
