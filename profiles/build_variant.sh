#!/bin/bash
# usage: build_variant.sh <tag> "<extra nvcc flags>"   -> build_variants/lib_<tag>.so
cd "$(dirname "$0")/.."; mkdir -p build_variants
make -s -C kvarq_b200/csrc OUT=$PWD/build_variants/lib_$1.so OBJ=$PWD/build_variants/obj_$1 EXTRA="$2" 2>&1 | grep -E "error" ; grep -E "registers|spill" build_variants/obj_$1/ptxas_scan_kernel.txt | head -2
