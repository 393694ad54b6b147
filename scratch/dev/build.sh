#!/bin/bash
# dev build: scratch/dev/build.sh <tag> [extra nvcc flags]  ->  scratch/dev/libfrx_<tag>.so (level-5 instantiations only)
set -e
cd /root/repo/flac_raster_b200/csrc
tag=$1; shift
out=/root/repo/scratch/dev
NV="/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -Xptxas -v --expt-relaxed-constexpr -DFRX_DEV_L5 $@"
$NV -c frx_encode.cu -o $out/enc_$tag.o 2> $out/enc_$tag.log &
[ -f $out/api.o ] && [ $out/api.o -nt frx_api.cu ] || $NV -c frx_api.cu -o $out/api.o 2> $out/api.log &
[ -f $out/dec.o ] && [ $out/dec.o -nt frx_decode.cu ] || $NV -c frx_decode.cu -o $out/dec.o 2> $out/dec.log &
[ -f $out/ing.o ] && [ $out/ing.o -nt frx_ingest.cu ] || $NV -c frx_ingest.cu -o $out/ing.o 2> $out/ing.log &
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/libfrx_$tag.so $out/enc_$tag.o $out/api.o $out/dec.o $out/ing.o
grep -A2 "frx_encode_kernel" $out/enc_$tag.log | grep -E "Used|spill" | head -8
