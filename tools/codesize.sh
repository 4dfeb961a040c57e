#!/bin/bash
# Static SASS rows per function of the list-8 quad kernel (M=10, float) and of its out-of-line helpers.
# usage: tools/codesize.sh [depth]
set -e
D=$(mktemp -d); R=$(cd "$(dirname "$0")/.." && pwd)
(cd $D && cuobjdump -xelf all $R/polarcodes_b200/libpolar_b200.so >/dev/null)
CUB=$(ls $D/decode_scl8q*.cubin)
nvdisasm -gi -c $CUB > $D/all.sass
# split per .text section
awk -v d=$D '/^\/\/-+ \.text\./{n=$0; sub(/.*\.text\./,"",n); sub(/ .*/,"",n); f=d"/"n".sec"} f{print > f}' $D/all.sass
for f in $D/*.sec; do
  n=$(basename $f .sec)
  case $n in *scl8q_kernelILi10ELi8E*|*fork_slow*|*heap_combine*) ;; *) continue;; esac
  rows=$(grep -cE '^\s+/\*[0-9a-f]{4,}\*/' $f || true)
  echo "== $n: $rows rows, $((rows*16)) bytes"
  python $R/tools/sass_rows.py $f $R/polarcodes_b200/csrc/decode_scl8q.cu ${1:-2} | head -${2:-14}
done
rm -rf $D
