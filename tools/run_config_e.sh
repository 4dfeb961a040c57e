#!/bin/bash
# BASELINE.json configs[4] as written: SCDecoderMinsum N=32768 K=16384, Monte-Carlo towards FER 1e-6, sharded over
# the GPUs of one box, one process per GPU, ncclAllReduce of the error counters after every slice (the C++ driver).
#   tools/run_config_e.sh [gpus] [frames per point] [max frame errors]      -> gpurun_out/config_e_*
# Also checks that G ranks and one rank end with the same counters on a short run (Philox keyed by the frame id).
set -e
R=$(cd "$(dirname "$0")/.." && pwd)
G=${1:-8}; FRAMES=${2:-100000000}; MAXERR=${3:-100}
OUT=$R/gpurun_out; mkdir -p $OUT
EXE=$R/polarcodes_b200/host/polarcodes_driver
make -s -C $R/polarcodes_b200/host
run_ranks() {   # dir world extra-env...
  local dir=$1 world=$2; shift 2
  local pids=()
  for r in $(seq 0 $((world-1))); do
    ( cd $dir && env WORLD_SIZE=$world RANK=$r LOCAL_RANK=$r MASTER_ADDR=127.0.0.1 MASTER_PORT=29700 PCB_NCCL_ID_FILE=$dir/nccl_id "$@" $EXE > $dir/rank$r.out 2> $dir/rank$r.err ) &
    pids+=($!)
  done
  local rc=0
  for p in "${pids[@]}"; do wait $p || rc=1; done
  return $rc
}
# 1) counters of G ranks == counters of one rank (short run at a point with errors)
T=$(mktemp -d)
mkdir -p $T/one $T/many
printf '1\n32768\n16384\n400000\n2\n2\n1\n0\n2\n2\n1.0 1.5\n' > $T/one/conf.txt
cp $T/one/conf.txt $T/many/conf.txt
( cd $T/one && PCB_SLICE=131072 PCB_CSV=$T/one/out.csv $EXE > $T/one/out.txt )
run_ranks $T/many $G PCB_SLICE=131072 PCB_CSV=$T/many/out.csv
cut -d, -f1-4 $T/one/out.csv > $OUT/config_e_equal_1gpu.csv
cut -d, -f1-4 $T/many/out.csv > $OUT/config_e_equal_${G}gpu.csv
if cmp -s $OUT/config_e_equal_1gpu.csv $OUT/config_e_equal_${G}gpu.csv; then echo "counters equal for 1 and $G GPUs:"; cat $OUT/config_e_equal_${G}gpu.csv; else echo "COUNTERS DIFFER"; cat $OUT/config_e_equal_1gpu.csv $OUT/config_e_equal_${G}gpu.csv; exit 1; fi
# 2) the run itself
D=$T/run; mkdir -p $D
printf "1\n32768\n16384\n$FRAMES\n2\n2\n1\n0\n2\n7\n1.0 1.25 1.5 1.75 2.0 2.25 2.5\n" > $D/conf.txt
t0=$(date +%s.%N)
run_ranks $D $G PCB_MAX_ERRORS=$MAXERR PCB_CSV=$D/out.csv PCB_CKPT=$D/run.ckpt
t1=$(date +%s.%N)
cp $D/out.csv $OUT/config_e_${G}gpu.csv
cp $D/SC_1_32768_16384_L1_CRC0_*.txt $OUT/config_e_${G}gpu_result.txt
cp $D/rank0.out $OUT/config_e_${G}gpu_stdout.txt
echo "wall $(python3 -c "print(round($t1 - $t0, 1))") s on $G GPUs" | tee $OUT/config_e_${G}gpu_wall.txt
cat $OUT/config_e_${G}gpu.csv
