"""BASELINE configs[1] (Cornell box, path tracing, 1920x1080) in 4-spp steps: Mrays/s, kernel launches per step and the effect of the wave size and of
`adaptive_iterations` (the fixed launch train).  usage: python tools/cornell_bench.py [steps=20]  -> one JSON line per variant"""
import json
import sys
import time

sys.path.insert(0, "/root/repo")
import bench  # noqa: E402
from libyafaray_b200 import Context, scenes  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
y = scenes.cornell_box(1920, 1080, 256, path_samples=1, bounces=8, raydepth=8, rr_min_bounces=2)
desc, keep = y.scene_desc()
for wave, adaptive in ((1 << 23, 0), (1 << 23, 1), (1 << 21, 0), (1 << 21, 1)):
    ctx = Context(0)
    ctx.set_option("wave_size", wave); ctx.set_option("adaptive_iterations", adaptive)
    ctx.upload(desc, keep); ctx.film_clear()
    for i in range(3):
        ctx.render_pass(256, 0, sample_begin=4 * i, sample_end=4 * i + 4)
    ctx.sync(); ctx.reset_stats()
    t0 = time.perf_counter()
    for i in range(steps):
        ctx.render_pass(256, 0, sample_begin=12 + 4 * i, sample_end=16 + 4 * i)
    ctx.sync()
    t = time.perf_counter() - t0
    st = ctx.stats()
    print(json.dumps({"wave_size": wave, "adaptive_iterations": adaptive, "mrays_s": bench.total_rays(st) / t / 1e6, "ms_per_step": t / steps * 1e3,
                      "kernel_launches_per_step": st["kernel_launches"] / steps, "waves_per_step": st["waves"] / steps}), flush=True)
    ctx.close()
