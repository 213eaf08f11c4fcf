"""Development tool: run-to-run noise floor of the REFERENCE itself (schedule-dependent Russian roulette / one-light choice):
two multithreaded reference renders of each converged parity case at the same spp and sampling offset."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import parity_cases as pc, oracle_util as ou
from libyafaray_b200.render import film_to_image
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
out = {}
for name, (build, _) in pc.CONVERGED.items():
    imgs = []
    for thr in (8, 5):
        t = time.time()
        f, info = ou.run_reference_render(build(spp), threads=thr, flavour="fast", timeout=36000)
        imgs.append(film_to_image(f)[..., :3])
        print(name, spp, "threads", thr, "%.1fs" % (time.time() - t)); sys.stdout.flush()
    print(name, "ref-vs-ref rel_rmse", ou.rel_rmse(imgs[0], imgs[1])); sys.stdout.flush()
    out[name + "_a"], out[name + "_b"] = imgs
np.savez_compressed("/tmp/noise_floor_%d.npz" % spp, **out)
