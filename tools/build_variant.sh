#!/bin/bash
# tools/build_variant.sh NAME "<flags for srt_api.cu>" "<flags for srt_team_tu.cu>"  -> tools/_variants/NAME.so
set -e
cd "$(dirname "$0")/.."
NV="nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -fmad=false -lineinfo -Xcompiler -fPIC -Xcompiler -fno-fast-math -w"
mkdir -p tools/_variants /tmp/srtv_$1
$NV $2 -c -o /tmp/srtv_$1/a.o stanford_raytracer_b200/csrc/srt_api.cu &
$NV $3 -c -o /tmp/srtv_$1/t.o stanford_raytracer_b200/csrc/srt_team_tu.cu &
g++ -O2 -std=c++17 -fPIC -pthread -c -o /tmp/srtv_$1/i.o stanford_raytracer_b200/csrc/srt_io.cpp &
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o tools/_variants/$1.so /tmp/srtv_$1/a.o /tmp/srtv_$1/t.o /tmp/srtv_$1/i.o -lpthread
rm -rf /tmp/srtv_$1
