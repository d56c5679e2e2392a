#!/bin/bash
# development aid (GPU box): rebuild me_search.o with each -D variant and time the decision kernels
cd kvazaar_b200/csrc
for v in "$@"; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden -fmad=false --expt-relaxed-constexpr $v -c me_search.cu -o build/me_search.o &&
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o ../libkvz_cuda.so build/*.o &&
  (cd ../..; echo "== variant $v"; python tools/time_me.py 2>/dev/null | head -2)
done
