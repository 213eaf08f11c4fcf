#!/bin/bash
# ncu --set full of the traversal kernels on the 10^7-triangle bench scene (iterations 1 and 2 of the first wave: incoherent extension and
# shadow rays).  usage: tools/ncu_trace.sh <tag> [trace_probe args]   (YAFGPU_LIB selects a kernel variant)
set -u
TAG=${1:-r2}; shift || true
OUT=gpurun_out; mkdir -p $OUT
REP=/tmp/yg_trace_$TAG
python tools/trace_probe.py "$@" > $OUT/trace_probe_$TAG.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:'k_trace<|k_shadow<' -s 2 -c 4 -f -o $REP \
    python tools/trace_probe.py "$@" > $OUT/ncu_trace_$TAG.log 2>&1
ncu -i $REP.ncu-rep --page raw --csv > $OUT/ncu_trace_raw_$TAG.csv 2>/dev/null
python tools/ncu_summary.py $OUT/ncu_trace_raw_$TAG.csv > $OUT/ncu_trace_summary_$TAG.txt 2>&1
ncu -i $REP.ncu-rep --page source --csv --print-source cuda,sass --kernel-name-base demangled -k regex:'k_trace<' -c 1 2>/dev/null | head -c 20000000 > $OUT/ncu_trace_source_$TAG.csv
ls -la $OUT | tail -6
