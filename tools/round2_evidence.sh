#!/bin/bash
# Round-2 evidence run on one B200: the contract bench (both arms), its ncu launch list and one ncu --set full capture of the traversal
# and shading kernels on the bench scene.  Outputs land in gpurun_out/ and are copied to profiles/ by hand.
set -u
OUT=gpurun_out; mkdir -p $OUT
(time python bench.py --impl reference --steps ${STEPS:-8} --warmup 3) > $OUT/r2_bench_ref_n1.json 2> $OUT/r2_bench_ref_n1.err
(time python bench.py --steps ${STEPS:-8} --warmup 3) > $OUT/r2_bench_n1.json 2> $OUT/r2_bench_n1.err
tail -c 300 $OUT/r2_bench_n1.err
python bench.py --steps 2 --warmup 1 --no-cpu > $OUT/r2_bench_short.json 2>/dev/null && \
timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2500 --csv --log-file $OUT/r2_launches_ncu.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu > $OUT/r2_ncu_launches.log 2>&1
python tools/make_ncu_traffic.py $OUT/r2_launches_ncu.csv $OUT/r2_bench_short.json r2 > $OUT/r2_launches.md 2>&1
cp profiles/ncu_traffic.json $OUT/ncu_traffic.json; cp profiles/r2_launches.md $OUT/ 2>/dev/null
REP=/tmp/yg_r2_full
timeout 1500 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:'k_trace<|k_shadow<|k_shade<' -s 14 -c 14 -f -o $REP \
    python tools/trace_probe.py 1e7 2 > $OUT/r2_ncu_full.log 2>&1
ncu -i $REP.ncu-rep --page raw --csv > $OUT/r2_ncu_raw_full.csv 2>/dev/null
python tools/ncu_summary.py $OUT/r2_ncu_raw_full.csv > $OUT/r2_ncu_summary.txt 2>&1
ls -la $OUT | tail -12
