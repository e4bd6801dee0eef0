#!/bin/bash
# A/B builds of the JPEG encoder for tools/gpu_jpeg_ab.sh: tools/jpeg_selftest_v<name> (git-ignored, they travel with gpurun)
cd "$(dirname "$0")/.."
build() { nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo $2 -o tools/jpeg_selftest_v$1 tools/jpeg_selftest.cu && echo "built v$1 ($2)"; }
build base ""
build jdb4 "-DRCT_JDB=4"
build roll "-DRCT_JEMIT_UNROLL=0"
build roll128 "-DRCT_JEMIT_UNROLL=0 -DRCT_JENT=128"
build ent128 "-DRCT_JENT=128"
build all "-DRCT_JDB=4 -DRCT_JEMIT_UNROLL=0 -DRCT_JENT=128"
