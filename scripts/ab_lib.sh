#!/bin/bash
# Build a second library whose <file>.cu comes from git revision <rev>, for a same-box A/B through FEDREC_LIB:
#   bash scripts/ab_lib.sh HEAD gemm_sm100        ->  build/ab/libfedrec_ab.so  (all other objects from the current build)
#   FEDREC_LIB=$PWD/build/ab/libfedrec_ab.so python bench.py ...   vs   python bench.py ...
set -e
cd "$(dirname "$0")/.."
rev=$1; name=$2
mkdir -p build/ab
git show "$rev:csrc/$name.cu" > build/ab/$name.cu
for h in common.cuh sm100.cuh; do git show "$rev:csrc/$h" > build/ab/$h; done
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden \
     --expt-relaxed-constexpr -I include -I build/ab -c build/ab/$name.cu -o build/ab/$name.o
objs=$(ls build/obj/*.o | grep -v "/$name.o")
nvcc -shared -o build/ab/libfedrec_ab.so build/ab/$name.o $objs -gencode arch=compute_100a,code=sm_100a
ls -la build/ab/libfedrec_ab.so
