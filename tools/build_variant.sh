#!/bin/bash
# build_variant.sh NAME "EXTRA_FLAGS" [file-overrides dir]: compiles csrc/*.cu (with overrides) into lib_ab/lib_NAME.so
set -e
NAME=$1; EXTRA=$2; OVR=$3
ROOT=$(cd $(dirname $0)/.. && pwd); C=$ROOT/xhy-flash-attention_b200/csrc; OUT=$ROOT/xhy-flash-attention_b200/lib_ab
T=$(mktemp -d); mkdir -p $OUT
for f in $C/*.cu; do
  src=$f; [ -n "$OVR" ] && [ -f $OVR/$(basename $f) ] && src=$OVR/$(basename $f)
  ( nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --expt-relaxed-constexpr -Xcompiler -fPIC $EXTRA -I $ROOT/include -I $C -c $src -o $T/$(basename $f .cu).o ) &
done
wait
nvcc -shared -o $OUT/lib_$NAME.so $T/*.o -lcudart
rm -rf $T; ls -la $OUT/lib_$NAME.so
