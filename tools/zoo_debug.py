"""Development tool: isolates material x light combinations of the zoo scene against the live oracle."""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libyafaray_b200 import Context, render_scene, film_to_image
from tests import oracle_util as ou, parity_cases as pc

names = ["gloss", "coated", "rough", "aniso", "mirr", "glas", "sd_on", "blend"]
ctx = Context(0)
integ = sys.argv[1] if len(sys.argv) > 1 else "directlighting"
sel = sys.argv[2].split(',') if len(sys.argv) > 2 else ['sun', 'meshlight', 'ibl']
for lights in [dict(sun=(n == 'sun'), meshlight=(n == 'meshlight'), ibl=(n == 'ibl')) for n in sel]:
    for k in range(8):
        for ts in (False, True):
            y = pc.material_zoo_scene(96, 54, 2, integrator=integ, raydepth=3, only_mat=k, transp_shad=ts, rr_min_bounces=99, **lights)
            g = film_to_image(render_scene(y, ctx=ctx))
            rf, _ = ou.run_reference_render(y)
            r = film_to_image(rf)
            d = np.abs(g[..., :3] - r[..., :3])
            print(json.dumps({"lights": [n for n, v in lights.items() if v], "mat": names[k], "tr_shad": ts, "max": float(d.max()), "rmse": float(np.sqrt((d ** 2).mean())),
                              "mean_g": float(g[..., :3].mean()), "mean_r": float(r[..., :3].mean()), "nbad": int((d.max(axis=2) > 1e-4).sum())}))
            sys.stdout.flush()
