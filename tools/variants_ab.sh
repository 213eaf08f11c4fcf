#!/bin/bash
# default library vs every libyafaray_b200/variants/*.so on the 10^7-triangle mesh (development).  Output: gpurun_out/variants_ab.jsonl
mkdir -p gpurun_out; : > gpurun_out/variants_ab.jsonl
run() { line=$(timeout 600 python tools/mesh_bench.py 1e7 2 "${@:2}" 2>/dev/null | tail -1); echo "{\"args\": \"$*\", \"r\": $line}" >> gpurun_out/variants_ab.jsonl; }
unset YAFGPU_LIB; run default $ARGS
for so in libyafaray_b200/variants/*.so; do export YAFGPU_LIB=$PWD/$so; run $(basename $so .so) $ARGS; done
python - <<'PY'
import json
for l in open('gpurun_out/variants_ab.jsonl'):
    try: d=json.loads(l)
    except Exception: print("bad", l[:100]); continue
    r=d["r"]; rn=r["render"]
    print("%-40s"%d["args"], "build %.1f | closest %.0f shadow %.0f Mr/s | render %.0f Mr/s trace %.1f shadow %.1f shade %.1f total %.1f"%(r["bvh_build_ms"],
          r["incoherent_closest"]["mrays_s"], r["incoherent_shadow"]["mrays_s"], rn["mrays_s_device"], rn["ms_trace"], rn["ms_shadow"], rn["ms_shade"], rn["ms_total"]))
PY
