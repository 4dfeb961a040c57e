#!/bin/bash
# A/B of library variants on config A (SC min-sum N=1024, fp64 and fp32): tools/ab_sc.sh var1 var2 ...
for v in "$@"; do
  if [ $v = main ]; then p=""; else p=polarcodes_b200/build/var_$v.so; fi
  echo "$v: $(PCB_LIB_PATH=$p python tools/config_sweep.py 2>&1 | head -2 | sed -E 's/.*\) +([0-9.]+) Mframes.*/\1/' | tr '\n' ' ')"
done
