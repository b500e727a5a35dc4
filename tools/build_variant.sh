#!/bin/bash
# usage: tools/build_variant.sh NAME file.cu "-DFOO=1 ..."   -> telomeri_b200/build/variants/NAME.so
# (the other objects are taken from the regular build; for kernel tuning experiments with HERA_B200_LIB)
set -e
cd "$(dirname "$0")/.."
name=$1; src=$2; defs=$3
mkdir -p telomeri_b200/build/variants
obj=telomeri_b200/build/variants/$name.o
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --fmad=false -Xcudafe --diag_suppress=20011,--diag_suppress=20014 \
  -Xptxas -v $defs -c telomeri_b200/csrc/$src -o $obj 2>&1 | grep -A2 "${4:-paf_parse_kernel}" | grep -E "registers|spill" || true
others=$(ls telomeri_b200/build/*.o | grep -v "/${src%.cu}.o")
nvcc -shared -o telomeri_b200/build/variants/$name.so $obj $others -lcudart
echo built $name
