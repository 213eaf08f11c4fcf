"""Generates the golden fixtures under tests/golden/ by running the oracle (the reference's own CPU code, oracle/_ref).
Usage: python tools/make_golden.py [exact|converged|<case name> ...] [--threads N]
The converged references are rendered with the `fast` flavour (the reference's Release flags) on N threads."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import oracle_util as ou, parity_cases as pc  # noqa: E402


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    threads = 8
    for a in sys.argv[1:]:
        if a.startswith("--threads="):
            threads = int(a.split("=")[1])
    want = args or ["exact", "hits", "converged"]
    for name, build in list(pc.EXACT.items()) + list(pc.FRONTEND_ONLY.items()):
        if "exact" in want or name in want:
            t0 = time.time()
            film = ou.reference_film(build(), name, write=True)
            print(name, film.shape, "%.1fs" % (time.time() - t0))
    for name, build in pc.TRACE.items():
        if "hits" in want or name in want:
            lo, hi = pc.TRACE_BOUNDS[name]
            rays = pc.trace_rays(8192, lo, hi)
            h = ou.reference_hits(build(), rays, 0, key="hits_" + name, write=True)
            sr = rays.copy(); sr[:, 3] = 5e-4; sr[:, 7] = np.random.default_rng(5).uniform(0.1, 6.0, len(sr)).astype(np.float32)
            s = ou.reference_hits(build(), sr, 1, key="shadow_" + name, write=True)
            print(name, "hits", int((h["tri"] >= 0).sum()), "shadowed", int((s["tri"] >= 0).sum()))
    for name, (build, spp) in pc.CONVERGED.items():
        if "converged" in want or name in want:
            t0 = time.time()
            y = build(spp)
            film, info = ou.run_reference_render(y, threads=threads, flavour="fast", timeout=36000)
            w = film[..., 4:5]
            img = np.where(w != 0, film[..., :3] / np.where(w != 0, w, 1), 0).astype(np.float32)
            np.savez_compressed(os.path.join(ou.GOLDEN, name + "_converged.npz"), image=img, spp=spp, rendert_s=info.get("rendert_s", 0.0), threads=threads)
            print(name, img.shape, spp, "spp %.1fs" % (time.time() - t0), info)


if __name__ == "__main__":
    main()
