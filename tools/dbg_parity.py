"""Development aid: GPU vs the live oracle on variants of a scene, to bisect a parity difference."""
import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from libyafaray_b200 import Context, render_scene, film_to_image, scenes
from tests import oracle_util as ou
ctx = Context(0)
def run(**kw):
    spp = kw.pop("spp")
    y = scenes.cornell_box(96, 54, spp, **kw)
    ref, _ = ou.run_reference_render(y)
    g, r = film_to_image(render_scene(y, ctx=ctx)), film_to_image(ref)
    d = g[..., :3] - r[..., :3]
    flip = np.abs(d).max(axis=2) > 2e-5 * max(1.0, r[..., :3].max())
    ys, xs = np.nonzero(flip)
    print(kw, "flips", len(ys), "max", float(np.abs(d).max()), "bbox", (ys.min(), ys.max(), xs.min(), xs.max()) if len(ys) else None,
          "gpu>ref", int((d[flip].sum(axis=1) > 0).sum()), "gpu<ref", int((d[flip].sum(axis=1) < 0).sum()))
    for y_, x_ in list(zip(ys, xs))[:4]:
        print("    ", (y_, x_), "gpu", g[y_, x_, :3], "ref", r[y_, x_, :3])
for rd in (1, 2, 3):
    run(spp=1, integrator="directlighting", dispersion=30.0, rough_glass=True, raydepth=rd)
run(spp=1, integrator="directlighting", dispersion=0.0, rough_glass=True, raydepth=3)
