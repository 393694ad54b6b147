#!/bin/bash
# dev build: scratch/dev/build.sh <tag> [extra nvcc flags]  ->  scratch/dev/libfrx_<tag>.so
# only frx_encode.cu is recompiled (level-5 instantiations, ~20 s); the other objects come from the production build in csrc/
set -e
cd /root/repo/flac_raster_b200/csrc
tag=$1; shift
out=/root/repo/scratch/dev
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -Xptxas -v --expt-relaxed-constexpr -DFRX_DEV_L5 "$@" -c frx_encode.cu -o $out/enc_$tag.o 2> $out/enc_$tag.log
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/libfrx_$tag.so $out/enc_$tag.o frx_api.o frx_decode.o frx_decode_w22.o frx_ingest.o
grep -A2 "frx_encode_kernel" $out/enc_$tag.log | grep -E "Used|spill" | head -4
