import json, sys, time
sys.path.insert(0, "/root/repo")
import bench
from libyafaray_b200 import Context, scenes
y = scenes.cornell_box(1920, 1080, 256, path_samples=1, bounces=8, raydepth=8, rr_min_bounces=2)
desc, keep = y.scene_desc()
for opts in ({}, {"ray_regeneration": 0}, {"ray_regeneration": 2}, {"ray_regeneration": 3}, {"regen_threshold": 28}, {"regen_threshold": 12}, {"shade_streams": 1}):
    ctx = Context(0, 2)
    for k, v in opts.items(): ctx.set_option(k, v)
    ctx.upload(desc, keep); ctx.film_clear()
    for i in range(3): ctx.render_pass(256, 0, sample_begin=4 * i, sample_end=4 * i + 4)
    ctx.sync(); ctx.reset_stats()
    t0 = time.perf_counter()
    for i in range(20): ctx.render_pass(256, 0, sample_begin=12 + 4 * i, sample_end=16 + 4 * i)
    ctx.sync(); t = time.perf_counter() - t0
    st = ctx.stats()
    print(json.dumps({"opts": opts, "mrays_s": round(bench.total_rays(st) / t / 1e6), "ms_per_step": round(t / 20 * 1e3, 2), "ms_trace": round(st["ms_trace"] / 20, 2), "ms_shadow": round(st["ms_shadow"] / 20, 2), "ms_shade": round(st["ms_shade"] / 20, 2)}), flush=True)
    ctx.close()
