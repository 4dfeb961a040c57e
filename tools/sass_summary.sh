#!/bin/bash
# SASS summary of the tuned kernels of the built library (committed as profiles/r2_sass_summary.txt): rows and bytes of the
# instantiations the BASELINE configurations run, their opcode histogram, and the mnemonics that show what the kernels are
# built from (FMNMX.XORSIGN / FADD2 / CREDUX / LDGSTS in the list kernel; DSETP with |.| modifiers, LDGSTS with an L2
# policy in the SC kernel; PRMT / VIMNMX.S16x2 in the 8-bit kernel).  No tensor-core or TMA mnemonic is expected: the path is a
# dependent tree walk (north_star).   usage: tools/sass_summary.sh > profiles/r2_sass_summary.txt
set -e
R=$(cd "$(dirname "$0")/.." && pwd); D=$(mktemp -d)
(cd $D && cuobjdump -xelf all $R/polarcodes_b200/libpolar_b200.so >/dev/null)
for spec in "decode_scl8q:scl8q_kernelILi10ELi8EfLi0E:SCL-8 N=1024 float batch (headline C)" "decode_scl8q:scl8q_kernelILi10ELi8EfLi1E:SCL-8 N=1024 double batch (narrowing prologue)" \
            "decode_scl8q:scl8q_kernelILi11ELi32EfLi0E:SCL-32 N=2048 (config D)" "decode_scl8q:scl8q_kernelILi11ELi4EfLi0E:SCL-4 N=2048 (conf.txt)" \
            "decode_scl8q:fork_slowILi8ELb0E:fork_slow<8> (out of line)" "decode_sct:sct_kernelIdLb0ELi4ELb1ELb0E:SC min-sum fp64, 4 warps per tile, global level slice, heap in shared memory (config A)" \
            "decode_sct:sct_kernelIdLb1ELi1ELb1ELb0E:exact SC, 1 warp per tile (config B)" "decode_sct:sct_kernelIdLb0ELi4ELb1ELb1E:SC min-sum fp64 n > 2048, levels and heap in the global slice (config E)" "decode_sc8:decode_sc8_kernel:8-bit SC" \
            "decode_sc32:sc32big_kernelIdLb0E:lane-per-frame SC min-sum fp64 n > 2048 (PCB_SC_LANE=1)"; do
  IFS=: read file pat label <<< "$spec"
  CUB=$(ls $D/${file}*.cubin | head -1)
  nvdisasm -c $CUB > $D/all.sass
  awk -v p="$pat" '/^\/\/-+ \.text\./{on = index($0, p) > 0} on' $D/all.sass > $D/k.sass
  rows=$(grep -cE '^\s+/\*[0-9a-f]{4,}\*/' $D/k.sass || true)
  echo "== $label [$pat]: $rows SASS rows, $((rows*16)) bytes"
  grep -oE '^\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P[0-9T]+ )?[A-Z0-9_.]+' $D/k.sass | awk '{print $NF}' | sed 's/\..*//' | sort | uniq -c | sort -rn | head -14 | awk '{printf "%s %s, ", $2, $1} END {print ""}'
  grep -oE '(FMNMX[.A-Z]*XORSIGN[.A-Z]*|FADD2|CREDUX[.A-Z0-9]*|LDGSTS[.A-Z0-9]*|DSETP[.A-Z]*|DMUL|DADD|PRMT|VIMNMX[.A-Z0-9]*|VIADD[.A-Z0-9]*|MUFU[.A-Z0-9]*|BAR[.A-Z]*|UTC[A-Z0-9.]*|UBLKCP[.A-Z0-9]*)' $D/k.sass | sort | uniq -c | sort -rn | head -10 | awk '{printf "%s x%s  ", $2, $1} END {print ""}'
done
rm -rf $D
