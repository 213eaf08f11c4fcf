"""Find a non-finite film sample of the bench workload: which pass (sample offset), which pixel, which sample index.
usage: python tools/nan_probe.py [passes=8] [spp_per_pass=128]"""
import sys

import numpy as np

sys.path.insert(0, "/root/repo")
import bench  # noqa: E402
from libyafaray_b200 import Context  # noqa: E402


class A:
    tris, width, height = bench.N_TRIS, bench.WIDTH, bench.HEIGHT


passes = int(sys.argv[1]) if len(sys.argv) > 1 else 8
S = int(sys.argv[2]) if len(sys.argv) > 2 else 128
y = bench.build_scene(A, S)
desc, keep = y.scene_desc()
ctx = Context(0); ctx.upload(desc, keep)
found = None
if "at" in sys.argv:   # ... at <px> <py> <pass offset> <sample>: that one sample only (for the -DYG_NAN_TRAP build, YAFGPU_LIB=...)
    a = sys.argv.index("at")
    px, py, off, smp = (int(v) for v in sys.argv[a + 1:a + 5])
    mask = np.zeros((A.height, A.width), np.uint8); mask[py, px] = 1
    ctx.film_clear()
    ctx.render_pass(S, off, adaptive=True, mask=mask, sample_begin=smp, sample_end=smp + 1)
    f = ctx.read_film()
    print("film at the pixel", f[py, px].tolist(), "non-finite pixels", int((~np.isfinite(f).all(axis=2)).sum()))
    sys.exit(0)
for i in range(passes):
    ctx.film_clear()
    ctx.render_pass(S, i * S)
    f = ctx.read_film()
    bad = ~np.isfinite(f).all(axis=2)
    print("pass", i, "offset", i * S, "non-finite pixels", int(bad.sum()), np.argwhere(bad)[:5].tolist(), flush=True)
    if bad.any() and found is None:
        found = (i, np.argwhere(bad)[0])
if found:
    i, (py, px) = found
    mask = np.zeros((A.height, A.width), np.uint8); mask[py, px] = 1
    for s in range(S):
        ctx.film_clear()
        ctx.render_pass(S, i * S, adaptive=True, mask=mask, sample_begin=s, sample_end=s + 1)
        f = ctx.read_film()
        if not np.isfinite(f).all():
            print("pixel", (int(px), int(py)), "pass offset", i * S, "sample", s, "film", f[~np.isfinite(f).all(axis=2)][:3].tolist(), flush=True)
