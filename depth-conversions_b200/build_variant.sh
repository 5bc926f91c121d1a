#!/bin/bash
# usage: build_variant.sh NAME "-DFLAG=.. ..." [only-TU ...]  -> variants/libdc_NAME.so (A/B kernel experiments)
# Builds the whole library with the extra -D switches into variants/NAME/ (objects are
# reused when only some translation units change) and leaves the .so next to it.
set -e
cd "$(dirname "$0")"
N=$1; F=$2
make -s -j${JOBS:-8} L=variants/$N DC_DEFS="$F" variants/$N/libdepthconv_b200.so
cp variants/$N/libdepthconv_b200.so variants/libdc_$N.so
cp variants/$N/ptxas.log variants/ptxas_$N.log
rm -rf variants/$N
