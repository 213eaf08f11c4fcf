#!/bin/bash
# 1/2/4/8-GPU runs of the contract bench only (quick refresh of profiles/r1_scale_bench.jsonl); usage on an 8-GPU box: tools/scale_bench_only.sh [tag]
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
NG=$(nvidia-smi -L | wc -l)
: > $OUT/scale_bench_$TAG.json
timeout 300 python bench.py --gpus 1 --steps 16 --warmup 3 --no-cpu 2>$OUT/scale_$TAG.log | tail -1 >> $OUT/scale_bench_$TAG.json
for N in 2 4 8; do
  [ $N -le $NG ] || continue
  timeout 300 $TR --nproc-per-node $N --master-port $((29500 + N)) bench.py --gpus $N --steps 16 --warmup 3 --no-cpu 2>>$OUT/scale_$TAG.log | grep '^{' | tail -1 >> $OUT/scale_bench_$TAG.json
done
python - <<PY
import json
for l in open("$OUT/scale_bench_$TAG.json"):
    l = l.strip()
    if l:
        d = json.loads(l); print("bench", d["n_gpus"], round(d["value"], 1), "Mrays/s", round(d["ms_per_step"], 2), "ms/step", "e2e", round(d["e2e"]["value"], 1))
PY
