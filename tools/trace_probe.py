"""Smallest run that shows the traversal kernels on the bench scene (for ncu): upload the N-triangle mesh, render one pass of `spp`
samples per pixel.  usage: python tools/trace_probe.py [tris] [spp] [leaf=N]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libyafaray_b200 import scenes, Context  # noqa: E402

n_tris = int(float(sys.argv[1])) if len(sys.argv) > 1 else 10_000_000
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 2
y = scenes.procedural_mesh(n_tris, 1920, 1080, 16)
desc, keep = y.scene_desc()
ctx = Context(0)
for a in sys.argv:
    if a.startswith("leaf="):
        ctx.set_option("leaf_size", int(a[5:]))
    if a.startswith("opt:"):
        ctx.set_option(a[4:].split("=")[0], int(a.split("=")[1]))
ctx.upload(desc, keep)
ctx.film_clear()
ctx.render_pass(16, 0, sample_begin=0, sample_end=spp)
st = ctx.stats()
print({k: st[k] for k in ("rays_camera", "rays_extension", "rays_shadow", "bvh_nodes", "bvh_width", "ms_total")})
ctx.close()
