#!/bin/sh
# build_variant.sh <name> <extra nvcc flags...>: builds vulkan-pathtracer_b200/variants/libb200pt_<name>.so
# (A/B kernel experiments; selected at run time with B200PT_LIB=<path>)
set -e
cd "$(dirname "$0")/../vulkan-pathtracer_b200/csrc"
NAME="$1"; shift
OBJ=_obj_$NAME; mkdir -p $OBJ ../variants
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --fmad=false -Xcompiler -fPIC,-ffp-contract=off $*"
for f in api render noalloc bvh_build; do $NV -c $f.cu -o $OBJ/$f.o -Xptxas -v 2> $OBJ/$f.log & done
wait
g++ -O2 -std=c++17 -fPIC -ffp-contract=off -c host.cpp -o $OBJ/host.o
g++ -O2 -std=c++17 -fPIC -ffp-contract=off -I/usr/local/cuda/include $* -c comm.cpp -o $OBJ/comm.o
g++ -O2 -std=c++17 -fPIC -ffp-contract=off -c scene_load.cpp -o $OBJ/scene_load.o
g++ -O2 -std=c++17 -fPIC -ffp-contract=off -c image_decode.cpp -o $OBJ/image_decode.o
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libb200pt_$NAME.so $OBJ/*.o -lz -ldl -Xlinker --no-undefined
grep -E "k_shade|k_trace_persistent" -A3 $OBJ/render.log | grep -E "Used|spill" | sed "s/ptxas info    ://"
