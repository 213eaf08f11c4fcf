"""Generates the reference answers for the hit-id parity tests at scale (tests/test_gpu_scale.py) with the reference's own
Scene::intersect (oracle/_ref/yaf_ref_ieee hits):

  oracle/_ref/cache/hits_<cfg>_<digest>.npy   full answer (2^21 rays; git-ignored, travels to the GPU box with the snapshot)
  tests/golden/scale_hits_<cfg>.npz           committed subset: every 128th ray's hit record + the digest of the full ray set

  tests/golden/scale_dl_c4.npz                the reference film of a DirectLight render of the 10^6-triangle config-4 geometry
                                              (blend materials, 3 mesh lights, area light, sun) at 160x90, 1 spp

    python tools/make_scale_hits.py c4 c3
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import ray_sets as rs, oracle_util as ou  # noqa: E402


def main():
    for name in sys.argv[1:] or ["c4", "c3"]:
        t0 = time.time()
        y = rs.scale_scene(name)
        rays = rs.scale_rays(name, y)
        print(name, "scene + rays", round(time.time() - t0, 1), "s; digest", rs.rays_digest(rays), flush=True)
        t0 = time.time()
        h, how = rs.scale_reference_hits(name, y, rays)
        print(name, how, round(time.time() - t0, 1), "s; hit fraction", float((h["tri"] >= 0).mean()), flush=True)
        sub = np.arange(0, len(rays), 128)
        np.savez_compressed(os.path.join(ou.GOLDEN, "scale_hits_%s.npz" % name), index=sub.astype(np.int32),
                            hits=h[sub].view(np.uint8).reshape(len(sub), 16), digest=np.frombuffer(rs.rays_digest(rays).encode(), np.uint8))


    if not sys.argv[1:] or "c4" in sys.argv[1:]:
        t0 = time.time()
        film = ou.reference_film(rs.scale_scene("c4"), "scale_dl_c4", write=True)
        print("scale_dl_c4", film.shape, round(time.time() - t0, 1), "s", flush=True)


if __name__ == "__main__":
    main()
