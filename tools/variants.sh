#!/bin/bash
# Tuning builds of libseqalign_b200.so: tools/variants.sh <name> [-D...]  ->  sequencealigner_b200/_variants/<name>.so
# Select at run time with SAB_LIB=<path> (sequencealigner_b200/capi.py).
set -e
cd "$(dirname "$0")/../sequencealigner_b200/csrc"
name="$1"; shift
mkdir -p ../_variants
nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo \
  -Xcompiler -fPIC,-O3,-Wall,-Wno-unused-function -shared -cudart shared "$@" \
  -o ../_variants/$name.so sab_capi.cu sab_filter.cu sab_tables.cpp sab_h5.cpp -lz
echo "built _variants/$name.so"
