#!/bin/bash
# ncu --set full + source page of the largest shading kernel of the bench scene (k_shade<2, 0>: glossy, integrate tasks; second launch = iteration 1 of the first wave)
set -u
OUT=gpurun_out; mkdir -p $OUT
REP=/tmp/yg_shade
K=${1:-'k_shade<.int.2, .int.0>'}
timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"$K" -s 1 -c 1 -f -o $REP \
    python tools/trace_probe.py 1e7 2 > $OUT/ncu_shade.log 2>&1
ncu -i $REP.ncu-rep --page raw --csv > $OUT/ncu_shade_raw.csv 2>/dev/null
python tools/ncu_summary.py $OUT/ncu_shade_raw.csv > $OUT/ncu_shade_summary.txt 2>&1
ncu -i $REP.ncu-rep --page source --csv --print-source cuda,sass --kernel-name-base demangled 2>/dev/null | head -c 30000000 > $OUT/ncu_shade_source.csv
python tools/ncu_source_lines.py $OUT/ncu_shade_source.csv > $OUT/ncu_shade_source_lines.txt 2>&1
rm -f $OUT/ncu_shade_source.csv
head -60 $OUT/ncu_shade_source_lines.txt
