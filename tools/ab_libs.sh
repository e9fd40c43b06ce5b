#!/bin/bash
# Same-box A/B of two builds of the library: HEAD's sources against the working tree's.
#   here:       tools/ab_libs.sh build            -> build/ab_old.so (git HEAD), build/ab_new.so (working tree)
#   on the box: tools/ab_libs.sh run "cfg3 cfg2" ["old new x .."]  -> r2_ab numbers for build/ab_<v>.so in that order
cd "$(dirname "$0")/.."
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC"
if [ "$1" = build ]; then
  mkdir -p build/ab_src && rm -rf build/ab_src/* && git archive HEAD quorum_b200/csrc include | tar -x -C build/ab_src
  nvcc $FLAGS -o build/ab_old.so build/ab_src/quorum_b200/csrc/quorum_b200.cu && nvcc $FLAGS -o build/ab_new.so quorum_b200/csrc/quorum_b200.cu && echo built
else
  cp quorum_b200/libquorum_b200.so build/ab_keep.so
  for v in ${3:-old new old new}; do
    cp build/ab_$v.so quorum_b200/libquorum_b200.so
    echo "#### $v"; bash tools/r2_ab.sh "$2" "token" | grep -v "^    [a-z_]*offsets"
  done
  cp build/ab_keep.so quorum_b200/libquorum_b200.so
fi
