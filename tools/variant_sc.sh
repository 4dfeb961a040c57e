#!/bin/bash
# like tools/variant.sh, for decode_sc32.cu (or SRC=<another .cu of csrc/>)
set -e
R=$(cd "$(dirname "$0")/.." && pwd); N=$1; shift
B=$R/polarcodes_b200/build
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC -Xcompiler -ffp-contract=off \
  "$@" -I $R/polarcodes_b200/csrc -I $R/include -c ${SRC:-$R/polarcodes_b200/csrc/decode_sc32.cu} -o $B/var_$N.o
S=${SRC:-$R/polarcodes_b200/csrc/decode_sc32.cu}
OBJS=$(ls $B/*.o | grep -v "var_\|$(basename $S).o")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $B/var_$N.so $OBJS $B/var_$N.o -lcudart -ldl
echo $B/var_$N.so
