#!/bin/bash
# Experiment helper: build a variant of libawfl_b200.so whose ord-5 kernels are compiled with extra -D flags.
#   tools/build_variant.sh NAME "-DAWFL_MINB5=5 ..."  ->  pam_b200/variants/libawfl_NAME.so  (select with AWFL_B200_LIB=path)
set -e
cd "$(dirname "$0")/../pam_b200/csrc"
name=$1; flags=$2
mkdir -p ../variants ../build
nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xptxas -v --compiler-options -fPIC -DAWFL_ORD_INST=5 $flags \
  -c awfl_ord.cu -o ../build/var_${name}_ord5.o > ../build/var_${name}.ptxas.log 2>&1 || { grep -A3 error ../build/var_${name}.ptxas.log; exit 1; }
nvcc -gencode arch=compute_100a,code=sm_100a -shared ../build/awfl_b200.o ../build/pam_modules.o ../build/pam_interface_b200.o ../build/awfl_ord3.o ../build/var_${name}_ord5.o ../build/awfl_ord7.o ../build/awfl_ord9.o -o ../variants/libawfl_${name}.so -ldl
echo "== $name: $flags"
python ../../tools/ptxas_summary.py ../build/var_${name}.ptxas.log | grep -E "march_(state<5, [02], [01], [01]>|tracer<5, [02], [01]>)"
