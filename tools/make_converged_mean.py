"""Converged REFERENCE image the way BASELINE.md §3.6 prescribes: the mean of K >= 16 renders of the reference's own CPU integrator
(oracle/_ref/yaf_ref_ieee) with distinct `adv_base_sampling_offset` (/root/reference/src/common/environment.cc:773,807), each at
S samples per pixel.  The offsets are k * S, so the K renders use consecutive ranges of the QMC sequences: their mean IS the K*S-spp
render, up to the Russian-roulette stream.

Writes tests/golden/<name>_mean_converged.npz: image (float32 mean), spp_total, K, S, and the reference's own noise floor —
`loo_rmse`: relative RMSE of every single S-spp render against the mean of the OTHER K-1 (what one reference realisation at S spp
scores against a converged image), from which the floor at K*S spp follows as loo_rmse.mean() / sqrt(K) for independent renders.

    python tools/make_converged_mean.py cornell_pt 32 18 65536 16"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("LIBYAFARAY_B200_HOST_ONLY", "1")
from libyafaray_b200 import scenes, film_to_image  # noqa: E402
from tests import oracle_util as ou  # noqa: E402

CASES = {
    "cornell_pt": lambda w, h, spp: scenes.cornell_box(w, h, spp, path_samples=1, bounces=8, raydepth=8, rr_min_bounces=2),
    "cornell_pt_noglass": lambda w, h, spp: scenes.cornell_box(w, h, spp, glass=False, path_samples=1, bounces=4, raydepth=4, rr_min_bounces=1),
}


def main():
    name, w, h, S, K = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
    threads = int(sys.argv[6]) if len(sys.argv) > 6 else -1
    imgs = []
    t0 = time.time()
    for k in range(K):
        y = CASES[name](w, h, S)
        y.render_params["adv_base_sampling_offset"] = ("int", k * S)
        film, info = ou.run_reference_render(y, threads=threads, flavour="ieee", timeout=24 * 3600)
        imgs.append(film_to_image(film)[..., :3].astype(np.float64))
        print("render %d/%d  %.0f s total, rendert %.1f s" % (k + 1, K, time.time() - t0, info["rendert_s"]), flush=True)
    imgs = np.stack(imgs)
    mean = imgs.mean(axis=0)
    loo = []
    for k in range(K):
        others = (imgs.sum(axis=0) - imgs[k]) / (K - 1)
        loo.append(ou.rel_rmse(imgs[k], others))
    half = ou.rel_rmse(imgs[: K // 2].mean(axis=0), imgs[K // 2:].mean(axis=0))
    out = os.path.join(ou.GOLDEN, "%s_mean_converged.npz" % name)
    np.savez_compressed(out, image=mean.astype(np.float32), spp_total=K * S, K=K, S=S, width=w, height=h, loo_rmse=np.array(loo), half_vs_half_rmse=half,
                        seconds=time.time() - t0)
    print(out, "leave-one-out rel. RMSE of one %d-spp reference render: mean %.3e; two halves (%d spp each) against each other: %.3e"
          % (S, float(np.mean(loo)), K // 2 * S, half))


if __name__ == "__main__":
    main()
