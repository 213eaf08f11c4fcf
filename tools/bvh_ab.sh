#!/bin/bash
# Wide-BVH A/B (development): node width (default library = 8-wide, variants/libyafgpu_w4.so = 4-wide) x leaf size, on the 10^7-triangle
# mesh: incoherent closest/any-hit rates, wavefront render stage times, node / triangle visits per ray.  Output: gpurun_out/bvh_ab.jsonl
mkdir -p gpurun_out
: > gpurun_out/bvh_ab.jsonl
for lib in default w8; do
  if [ $lib = default ]; then unset YAFGPU_LIB; else export YAFGPU_LIB=$PWD/libyafaray_b200/variants/libyafgpu_$lib.so; fi
  for leaf in ${LEAVES:-1 2 4}; do
    for mode in time count; do
      extra=""; [ $mode = count ] && extra="count"
      line=$(timeout 600 python tools/mesh_bench.py 1e7 2 leaf=$leaf $extra 2>/dev/null | tail -1)
      echo "{\"lib\": \"$lib\", \"leaf\": $leaf, \"mode\": \"$mode\", \"r\": $line}" >> gpurun_out/bvh_ab.jsonl
    done
  done
done
python - <<'PY'
import json
for l in open('gpurun_out/bvh_ab.jsonl'):
    try: d=json.loads(l)
    except Exception as e: print("bad line", l[:200]); continue
    r=d["r"]; rn=r["render"]
    print(d["lib"], "leaf", d["leaf"], d["mode"], "build %.1f ms nodes %d (%.0f MB)"%(r["bvh_build_ms"], r["bvh_nodes"], r["bvh_mb"]),
          "| closest %.0f Mr/s n/r %s t/r %s | shadow %.0f Mr/s | render %.0f Mr/s trace %.1f shadow %.1f shade %.1f total %.1f ms"%(
          r["incoherent_closest"]["mrays_s"], r["incoherent_closest"]["nodes_per_ray"], r["incoherent_closest"]["tris_per_ray"], r["incoherent_shadow"]["mrays_s"],
          rn["mrays_s_device"], rn["ms_trace"], rn["ms_shadow"], rn["ms_shade"], rn["ms_total"]))
PY
