"""Load balance of the N-way image partition on the bench workload, measured on ONE GPU: partition k of N rendered alone, device time per partition.
The slowest partition ends a multi-GPU step, so max / mean is the efficiency the partition itself costs.  usage: python tools/partition_balance.py [N=8] [tile_size...]"""
import json
import sys
import time

import numpy as np

sys.path.insert(0, "/root/repo")
import bench  # noqa: E402
from libyafaray_b200 import Context  # noqa: E402


class A:
    tris, width, height = bench.N_TRIS, bench.WIDTH, bench.HEIGHT


N = int(sys.argv[1]) if len(sys.argv) > 1 else 8
y = bench.build_scene(A, 16)
desc, keep = y.scene_desc()
for ts in [int(a) for a in sys.argv[2:]] or [32]:
    desc.film.tile_size = ts
    ctx = Context(0)
    ctx.upload(desc, keep); ctx.film_clear()
    ctx.render_pass(16, 0, part_index=0, part_count=N); ctx.sync()
    ms = []
    for k in range(N):
        best = 1e9
        for rep in range(3):   # the same samples three times: the minimum is the partition's cost, the rest is host jitter
            t0 = time.perf_counter()
            ctx.render_pass(16, 16, part_index=k, part_count=N); ctx.sync()
            best = min(best, (time.perf_counter() - t0) * 1e3)
        ms.append(best)
    ms = np.array(ms)
    print(json.dumps({"parts": N, "tile_size": ts, "ms": [round(float(v), 2) for v in ms], "max_over_mean": round(float(ms.max() / ms.mean()), 4)}), flush=True)
    ctx.close()
