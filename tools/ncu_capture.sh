#!/bin/bash
# One `ncu --set full` capture of the hot kernels of the bench workload; exports CSV pages small enough to travel back in gpurun_out/.
# usage: tools/ncu_capture.sh <tag> [bench args...]
set -u
TAG=${1:-r1}; shift || true
OUT=gpurun_out
mkdir -p $OUT
REP=/tmp/yg_$TAG
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:'k_shade<\(int\)[0-9], \(int\)[01]>|k_trace<|k_shadow<' -s 4 -c 8 -f -o $REP \
    python bench.py --steps 1 --warmup 0 --no-cpu "$@" > $OUT/ncu_full_$TAG.log 2>&1
ncu -i $REP.ncu-rep --page raw --csv > $OUT/ncu_raw_$TAG.csv 2>/dev/null
ncu -i $REP.ncu-rep --page details --csv > $OUT/ncu_details_$TAG.csv 2>/dev/null
ncu -i $REP.ncu-rep --page source --csv --print-source cuda,sass --kernel-name-base demangled -k regex:'k_shade<\(int\)9, \(int\)0>' -c 1 2>/dev/null | head -c 30000000 > $OUT/ncu_source_shade_$TAG.csv
python tools/ncu_source_lines.py $OUT/ncu_source_shade_$TAG.csv 60 > $OUT/ncu_source_lines_$TAG.txt 2>&1
rm -f $OUT/ncu_source_shade_$TAG.csv
ls -la $OUT | tail -8
