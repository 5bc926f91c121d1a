#!/bin/bash
# usage: build_variant.sh NAME "-DFLAG=.. ..."   -> variants/libdc_NAME.so (A/B kernel experiments)
set -e
cd "$(dirname "$0")"
N=$1; F=$2
mkdir -p variants
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true -ftz=false -Xcompiler -fPIC,-ffp-contract=off,-Wall -Xptxas -v $F -c csrc/dc_api.cu -o variants/dc_api_$N.o 2> variants/ptxas_$N.log
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o variants/libdc_$N.so variants/dc_api_$N.o lib/dc_tables.o lib/dc_f64.o -lpthread
rm -f variants/dc_api_$N.o
