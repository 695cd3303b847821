#!/bin/bash
# Experiment helper: build a variant of libawfl_b200.so whose kernels of ONE order are compiled with extra -D flags.
#   tools/build_variant.sh NAME "-DAWFL_MINB5=5 ..." [ORD=5]  ->  pam_b200/variants/libawfl_NAME.so  (select with AWFL_B200_LIB=path)
set -e
cd "$(dirname "$0")/../pam_b200/csrc"
name=$1; flags=$2; ord=${3:-5}
mkdir -p ../variants ../build
nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xptxas -v --compiler-options -fPIC -DAWFL_ORD_INST=$ord $flags \
  -c awfl_ord.cu -o ../build/var_${name}_ord${ord}.o > ../build/var_${name}.ptxas.log 2>&1 || { grep -A3 error ../build/var_${name}.ptxas.log; exit 1; }
objs=""
for o in 3 5 7 9; do if [ $o = $ord ]; then objs="$objs ../build/var_${name}_ord${ord}.o"; else objs="$objs ../build/awfl_ord$o.o"; fi; done
nvcc -gencode arch=compute_100a,code=sm_100a -shared ../build/awfl_b200.o ../build/pam_modules.o ../build/pam_interface_b200.o $objs -o ../variants/libawfl_${name}.so -ldl
echo "== $name (ord $ord): $flags"
python ../../tools/ptxas_summary.py ../build/var_${name}.ptxas.log | grep -E "march_(state<$ord, [02], [01], [01]>|tracer<$ord, [02], [01]>)"
