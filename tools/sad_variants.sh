#!/bin/bash
# development aid (GPU box): rebuild ime_sad.o with each -D variant and time the box SAD stage
cd kvazaar_b200/csrc
for v in "$@"; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden -fmad=false --expt-relaxed-constexpr $v -c ime_sad.cu -o build/ime_sad.o &&
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o ../libkvz_cuda.so build/*.o &&
  (cd ../..; echo "== variant $v"; KVZ_EXP_ALLOC16=1 python tools/time_sad_box.py 2>&1 | tail -2)
done
