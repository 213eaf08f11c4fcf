#!/bin/bash
# 1/2/4/8-GPU runs of the contract bench (configs[1]) and of configs 3-5 (tools/config_bench.py) on one box; JSON lines into gpurun_out/scale_*.json
# usage (on an 8-GPU box): tools/scale_run.sh [tag]
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
NG=$(nvidia-smi -L | wc -l)
echo "gpus: $NG" > $OUT/scale_$TAG.log
: > $OUT/scale_bench_$TAG.json
: > $OUT/scale_configs_$TAG.json
timeout 300 python bench.py --gpus 1 --steps 16 --warmup 3 --no-cpu 2>>$OUT/scale_$TAG.log | tail -1 >> $OUT/scale_bench_$TAG.json
for N in 2 4 8; do
  [ $N -le $NG ] || continue
  timeout 300 $TR --nproc-per-node $N --master-port $((29500 + N)) bench.py --gpus $N --steps 16 --warmup 3 --no-cpu 2>>$OUT/scale_$TAG.log | grep '^{' | tail -1 >> $OUT/scale_bench_$TAG.json
done
timeout 300 python tools/config_bench.py c3 4 2 2>>$OUT/scale_$TAG.log | grep '^{' >> $OUT/scale_configs_$TAG.json
timeout 300 python tools/config_bench.py c4 4 2 2>>$OUT/scale_$TAG.log | grep '^{' >> $OUT/scale_configs_$TAG.json
for N in 2 4; do
  [ $N -le $NG ] || continue
  timeout 400 $TR --nproc-per-node $N --master-port $((29600 + N)) tools/config_bench.py c4 4 2 2>>$OUT/scale_$TAG.log | grep '^{' >> $OUT/scale_configs_$TAG.json
done
timeout 400 python tools/config_bench.py c5 3 1 2>>$OUT/scale_$TAG.log | grep '^{' >> $OUT/scale_configs_$TAG.json
if [ 8 -le $NG ]; then
  timeout 600 $TR --nproc-per-node 8 --master-port 29708 tools/config_bench.py c5 3 1 2>>$OUT/scale_$TAG.log | grep '^{' >> $OUT/scale_configs_$TAG.json
fi
tail -c 600 $OUT/scale_$TAG.log
python - <<PY
import json
for f in ("$OUT/scale_bench_$TAG.json", "$OUT/scale_configs_$TAG.json"):
    for l in open(f):
        l = l.strip()
        if not l: continue
        d = json.loads(l)
        print(d.get("config") if isinstance(d.get("config"), str) else "bench", d["n_gpus"], round(d.get("value", d.get("mrays_s", 0)), 1), "Mrays/s", round(d["ms_per_step"], 2), "ms/step")
PY
