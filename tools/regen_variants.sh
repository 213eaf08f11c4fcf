#!/bin/bash
# Development A/B: bench with every libyafaray_b200/variants/*.so (YAFGPU_LIB) next to the in-tree library
run() { python bench.py --steps 12 --warmup 3 --no-cpu 2>/dev/null | tail -1 | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('$1', round(j['value']), round(j['ms_per_step'],3), {k: round(v,2) for k,v in j['roofline']['stage_ms'].items()})"; }
run base
for so in libyafaray_b200/variants/*.so; do export YAFGPU_LIB=$PWD/$so; run $(basename $so .so); done
unset YAFGPU_LIB
run base_again
