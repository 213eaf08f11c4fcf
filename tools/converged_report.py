"""Prints rel. RMSE of the GPU render against each converged reference fixture (numbers quoted in DESIGN.md §7)."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libyafaray_b200 import render_scene, film_to_image, Context
from tests import oracle_util as ou, parity_cases as pc
ctx = Context(0)
for name, (build, _) in pc.CONVERGED.items():
    p = os.path.join(ou.GOLDEN, name + "_converged.npz")
    if not os.path.exists(p):
        continue
    z = np.load(p)
    img = film_to_image(render_scene(build(int(z["spp"])), ctx=ctx))[..., :3]
    np.savez_compressed(os.path.join(ROOT, "gpurun_out", "conv_%s.npz" % name), gpu=img, ref=z["image"])
    print(json.dumps({"case": name, "spp": int(z["spp"]), "rel_rmse": ou.rel_rmse(img, z["image"]), "mean_ratio": float(img.mean() / z["image"].mean())}))
