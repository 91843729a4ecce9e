#!/bin/bash
# SASS of the hot kernels as built in total-lagrangian-fea_b200/lib/libtlfea_b200.so (north_star: a committed listing per
# kernel).  Usage: tools/dump_sass.sh [tag]   -> profiles/<tag>_sass_<kernel>.txt + a mnemonic summary.
set -e
cd "$(dirname "$0")/.."
TAG=${1:-r2}
SO=total-lagrangian-fea_b200/lib/libtlfea_b200.so
dump() {  # $1 = output name, $2 = grep pattern on the mangled function name
  local fn
  fn=$(cuobjdump -sass "$SO" | grep -o "Function : [^ ]*" | awk '{print $3}' | grep -E "$2" | head -1)
  [ -z "$fn" ] && { echo "no kernel matches $2"; return; }
  cuobjdump -sass -fun "$fn" "$SO" 2>/dev/null > "profiles/${TAG}_sass_$1.txt"
  printf "%-28s %6d lines  UBLKCP %3d  LDGSTS %3d  DFMA %5d  LDS %4d  STS %4d  LDL %3d  STL %3d  (%s)\n" "$1" \
    "$(wc -l < profiles/${TAG}_sass_$1.txt)" "$(grep -c UBLKCP profiles/${TAG}_sass_$1.txt)" \
    "$(grep -c LDGSTS profiles/${TAG}_sass_$1.txt)" "$(grep -c DFMA profiles/${TAG}_sass_$1.txt)" \
    "$(grep -c 'LDS' profiles/${TAG}_sass_$1.txt)" "$(grep -c 'STS' profiles/${TAG}_sass_$1.txt)" \
    "$(grep -c 'LDL' profiles/${TAG}_sass_$1.txt)" "$(grep -c 'STL' profiles/${TAG}_sass_$1.txt)" "$fn"
}
{
dump pk_kernel_w4_svk        'pk_kernelILi4ELb0ELb1ELi0E'
dump pk_kernel_w4_svk_damp   'pk_kernelILi4ELb1ELb1ELi0E'
dump pk_kernel_w4_mr         'pk_kernelILi4ELb0ELb1ELi1E'
dump ek_kernel_2_8_svk       'ek_kernelILi2ELi8ELb0ELb1ELi0E'
dump ek_kernel_2_7_mr        'ek_kernelILi2ELi7ELb0ELb1ELi1E'
dump rk_kernel_10_5_3        'rk_kernelILi10ELi5ELi3E'
dump rk_kernel_8_5_3         'rk_kernelILi8ELi5ELi3E'
dump rk_kernel_16_3_3        'rk_kernelILi16ELi3ELi3E'
dump aek_3443_svk            'aek_kernelILi16ELi48ELi1ELb0ELb1ELi0E'
dump aek_3443_mr             'aek_kernelILi16ELi48ELi1ELb0ELb1ELi1E'
dump aek_3243_svk            'aek_kernelILi8ELi12ELi7ELb0ELb1ELi0E'
dump aek_3243_mr             'aek_kernelILi8ELi12ELi3ELb0ELb1ELi1E'
} | tee "profiles/${TAG}_sass_summary.txt"
