#!/bin/bash
# Ray-regeneration A/B on the 10^7-triangle mesh (development): schedule x refill threshold.  Output: gpurun_out/regen_ab.jsonl
mkdir -p gpurun_out; : > gpurun_out/regen_ab.jsonl
run() { line=$(timeout 600 python tools/mesh_bench.py 1e7 2 "$@" 2>/dev/null | tail -1); echo "{\"args\": \"$*\", \"r\": $line}" >> gpurun_out/regen_ab.jsonl; }
run leaf=2 count
run leaf=2
for thr in 12 20 28; do run leaf=2 opt:ray_regeneration=2 opt:regen_threshold=$thr; done
run leaf=2 opt:ray_regeneration=1 opt:regen_threshold=20
run leaf=2 opt:ray_regeneration=3 opt:regen_threshold=20
run leaf=4 opt:ray_regeneration=1 opt:regen_threshold=20
python - <<'PY'
import json
for l in open('gpurun_out/regen_ab.jsonl'):
    try: d=json.loads(l)
    except Exception: print("bad", l[:100]); continue
    r=d["r"]; rn=r["render"]
    print("%-55s"%d["args"], "build %.1f | closest %.0f shadow %.0f Mr/s | render %.0f Mr/s trace %.1f shadow %.1f shade %.1f total %.1f | n/r %s t/r %s"%(r["bvh_build_ms"],
          r["incoherent_closest"]["mrays_s"], r["incoherent_shadow"]["mrays_s"], rn["mrays_s_device"], rn["ms_trace"], rn["ms_shadow"], rn["ms_shade"], rn["ms_total"], rn.get("closest_nodes_per_ray"), rn.get("closest_tris_per_ray")))
PY
