#!/bin/sh
# Build experimental variants of libflsgpu.so into duckdb-fastlane_b200/exp/ (they travel to the GPU box):
#   profiles/exp.sh NAME "-DFLS_EXP=1" [NAME2 "-D..."] ...
# run there with  FLSGPU_LIB=$PWD/duckdb-fastlane_b200/exp/libflsgpu_NAME.so python profiles/colbench.py 916 5 l_comment
set -e
cd "$(dirname "$0")/../duckdb-fastlane_b200"
mkdir -p exp
while [ $# -ge 2 ]; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --fmad=false -Xcompiler -fPIC -shared -w \
       $2 csrc/flsgpu.cu csrc/fls_reader.cpp -o exp/libflsgpu_$1.so &
  shift 2
done
wait
