#!/bin/bash
# Builds librtseg from the committed sources (git HEAD) into gpurun_scratch/ab/libhead.so for in-box A/B runs:
#   RTSEG_LIB_PATH=/root/repo/gpurun_scratch/ab/libhead.so python bench.py ...   vs   python bench.py ...
set -e
cd "$(dirname "$0")/.."
rm -rf gpurun_scratch/ab && mkdir -p gpurun_scratch/ab/src
git archive HEAD realtime-semantic-segmentation-pytorch_b200/csrc include | tar -x -C gpurun_scratch/ab/src
cd gpurun_scratch/ab/src/realtime-semantic-segmentation-pytorch_b200/csrc
objs=""
for f in *.cu; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=default --expt-relaxed-constexpr -I../../include -c $f -o ${f%.cu}.o 2>/dev/null &
  objs="$objs ${f%.cu}.o"
done
wait
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o ../../../libhead.so $objs
cd ../../.. && rm -rf src && ls -la libhead.so
