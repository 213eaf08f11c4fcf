#!/bin/bash
# Round-2 scaling evidence on one 8xB200 box: the contract bench at N = 1, 2, 4, 8 (torchrun, NCCL film reduce) and the in-process yafgpu_group at N = 8.
set -u
OUT=gpurun_out; mkdir -p $OUT
python bench.py --steps 6 --warmup 3 --no-cpu 2>$OUT/r2_bench_n1_8box.err | grep '^{' > $OUT/r2_bench_n1_8box.json
for N in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N --steps 6 --warmup 3 2>$OUT/r2_bench_n$N.err | grep '^{' > $OUT/r2_bench_n$N.json
done
python tools/group_bench.py 8 5 2 2>$OUT/r2_group_n8.err | grep '^{' > $OUT/r2_group_n8.json
python - <<'PY'
import json
for n in ("r2_bench_n1_8box", "r2_bench_n2", "r2_bench_n4", "r2_bench_n8", "r2_group_n8"):
    try:
        d = json.load(open("gpurun_out/%s.json" % n))
        print(n, round(d.get("value", d.get("value_mrays_s", 0))), round(d.get("ms_per_step", 0), 1), round((d.get("e2e") or {}).get("value", 0)), d.get("finite"), d.get("ms_last_film_reduce"))
    except Exception as e:
        print(n, "failed", e)
PY
