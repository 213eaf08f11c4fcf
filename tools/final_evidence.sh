set -u
mkdir -p gpurun_out
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/final_ref.json 2> gpurun_out/final_ref.err; echo "ref rc=$?"
timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_r1n.csv python bench.py --steps 1 --warmup 0 --no-cpu > gpurun_out/ncu_launch_r1n.log 2>&1; echo "ncu rc=$?"
tail -c 600 gpurun_out/final_bench.json; tail -c 400 gpurun_out/final_ref.json
wc -l gpurun_out/launches_r1n.csv
