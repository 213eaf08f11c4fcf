"""Development tool: GPU renders of the converged parity cases (two independent sampling offsets) -> gpurun_out/conv_<name>.npz"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libyafaray_b200 import Context, render_scene, film_to_image
from tests import parity_cases as pc, oracle_util as ou
ctx = Context(0)
for name in (sys.argv[1:] or list(pc.CONVERGED)):
    build, spp = pc.CONVERGED[name]
    z = np.load(os.path.join(ou.GOLDEN, name + "_converged.npz"))
    imgs = []
    for off in (0, 7000003):
        y = build(int(z["spp"]))
        y.render_params["adv_base_sampling_offset"] = ("int", off)
        t = time.time()
        imgs.append(film_to_image(render_scene(y, ctx=ctx))[..., :3])
        print(name, off, "%.2fs" % (time.time() - t), "rel_rmse vs ref", ou.rel_rmse(imgs[-1], z["image"]))
    print(name, "gpu-vs-gpu", ou.rel_rmse(imgs[0], imgs[1]))
    np.savez_compressed(os.path.join(ROOT, "gpurun_out", "conv_%s.npz" % name), a=imgs[0], b=imgs[1], ref=z["image"])
