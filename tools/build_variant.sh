#!/bin/bash
# tools/build_variant.sh NAME "<extra nvcc flags>"  ->  tools/variants/NAME.so  (A/B builds; RTB200_LIB=... selects one)
set -e
cd "$(dirname "$0")/../raytracing-gpu_b200"
name=$1; shift
out=../tools/variants; mkdir -p $out/obj_$name
for f in api build sort trace; do
  /usr/local/cuda/bin/nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false \
    -Xcompiler -fPIC,-ffp-contract=off,-fno-fast-math -I../include -Icsrc $@ -c csrc/$f.cu -o $out/obj_$name/$f.o &
done
wait
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $out/$name.so $out/obj_$name/*.o build/host/*.o -lz -lpthread
rm -rf $out/obj_$name
echo built $out/$name.so
