#!/bin/bash
# A/B of library variants on one box: tools/ab.sh [rounds] var1 var2 ...  ("main" = the in-tree library)
R=${1:-2}; shift
for r in $(seq $R); do for v in "$@"; do
  if [ $v = main ]; then p=""; else p=polarcodes_b200/build/var_$v.so; fi
  PCB_LIB_PATH=$p python bench.py --steps 5 --warmup 3 --kernel-only --frames ${FRAMES:-2097152} ${BENCH_ARGS:-} 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', round(d['frames_per_s']/1e6,3), round(d['roofline']['kernel_ms'],4))"
done; done
