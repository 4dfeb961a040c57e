#!/bin/bash
# Rebuilds the committed profile summaries of the list-8 kernel from the captures of the bench command:
#   gpurun_out/r1_final_launches.csv  (ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv ... python bench.py --steps 2 --warmup 1 --no-cpu)
#   gpurun_out/r1_scl8q_final.ncu-rep (ncu --set full --clock-control none --import-source on -k regex:decode_scl8q -c 1 ... same command)
set -e
cd "$(dirname "$0")/.."
L=gpurun_out/r1_final_launches.csv; R=gpurun_out/r1_scl8q_final.ncu-rep
python tools/launch_summary.py $L "launch list of: python bench.py --steps 2 --warmup 1 --no-cpu (262144 frames/step, e2e legs included; first 400 launches, ncu gpu__time_duration.sum --clock-control none, cold-cache serialised)" > profiles/r1_scl8q_launches_summary.txt
python - <<'PY' >> profiles/r1_scl8q_launches_summary.txt
import csv
rows=[r for r in csv.reader(open('gpurun_out/r1_final_launches.csv')) if len(r)>14 and r[0].isdigit()]
pre=max(float(r[14]) for r in rows if 'prepass' in r[4]); dec=max(float(r[14]) for r in rows if 'decode_scl8q' in r[4]); gen=min(float(r[14]) for r in rows if 'decode_generic' in r[4])
print(f"\none device-resident step of 262144 frames: scl8_prepass {pre/1e6:.3f} ms + decode_scl8q {dec/1e6:.3f} ms + decode_generic (empty redo list) {gen/1e6:.3f} ms\n-> share of decode_scl8q in the step: {100*dec/(pre+dec+gen):.1f} %")
PY
cp $L profiles/r1_scl8q_launches.csv
python tools/ncu_summary.py $R 262144 > profiles/r1_scl8q_ncu_summary.txt
python tools/ncu_funcs.py $R polarcodes_b200/libpolar_b200.so kernelILi10ELi8 polarcodes_b200/csrc/decode_scl8q.cu 3 65536 > profiles/r1_scl8q_ncu_functions.txt 2>&1
python tools/ncu_lines.py $R polarcodes_b200/libpolar_b200.so kernelILi10ELi8 40 > profiles/r1_scl8q_ncu_lines.txt 2>&1
python - <<'PY'
import json, re
t=open('profiles/r1_scl8q_ncu_summary.txt').read()
g=lambda k: float(re.search(k+r'\s+([0-9.]+) Gbyte', t).group(1))*1e9
json.dump({"kernel": "decode_scl8q_kernel<10, 8, float>", "frames_per_launch": 262144, "dram_bytes_read": g('dram__bytes_read.sum'), "dram_bytes_write": g('dram__bytes_write.sum'),
 "source": "profiles/r1_scl8q_ncu_summary.txt (ncu --set full --clock-control none, one launch of python bench.py --steps 2 --warmup 1 --no-cpu)"}, open('profiles/r1_scl8q_traffic.json','w'), indent=1)
PY
tail -3 profiles/r1_scl8q_launches_summary.txt; head -14 profiles/r1_scl8q_ncu_summary.txt; cat profiles/r1_scl8q_traffic.json
