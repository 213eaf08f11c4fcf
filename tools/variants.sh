#!/bin/bash
# Kernel-variant A/B run (development): every libyafaray_b200/variants/*.so through the exact-parity tests, the bench and the mesh bench.
mkdir -p gpurun_out
for so in libyafaray_b200/variants/*.so; do
  tag=$(basename $so .so)
  export YAFGPU_LIB=$PWD/$so
  echo "=== $tag"
  timeout 600 python -m pytest tests -m gpu -q -x -k "matches_reference_film" 2>&1 | tail -1
  timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('bench Mrays/s %.0f ms/step %.2f stage %s'%(d['value'], d['ms_per_step'], {k:round(v,1) for k,v in r['stage_ms'].items()}))"
  timeout 300 python tools/mesh_bench.py 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())['render']
print('mesh Mrays/s %.0f trace %.1f shadow %.1f shade %.1f total %.1f'%(d['mrays_s'], d['ms_trace'], d['ms_shadow'], d['ms_shade'], d['ms_total']))"
done
