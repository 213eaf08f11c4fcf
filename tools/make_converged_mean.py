"""Converged REFERENCE image the way BASELINE.md §3.6 prescribes: the mean of K >= 16 renders of the reference's own CPU integrator
(oracle/_ref/yaf_ref_ieee) with distinct `adv_base_sampling_offset` (/root/reference/src/common/environment.cc:773,807), each at
S samples per pixel.  The offsets are k * S, so the K renders use consecutive ranges of the QMC sequences: their mean IS the K*S-spp
render, up to the Russian-roulette stream.

Incremental: the fixture tests/golden/<name>_mean_converged.npz keeps `block_means` / `block_counts` (the first run's 16 renders are one
block, every later render its own), so the tool can be left running to add renders; `image` is the mean over all of them, `K`, `S`,
`spp_total` say how many, `half_vs_half_rmse` is the relative RMSE between the means of the two halves of the blocks (the reference's own
noise floor at HALF the total spp), `loo_rmse` the leave-one-out RMSE of the single-render blocks against the mean of everything else.

    python tools/make_converged_mean.py cornell_pt 32 18 65536 <renders to add> [threads]"""
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


def save(out, means, counts, S, w, h, seconds):
    means, counts = np.asarray(means, np.float64), np.asarray(counts, np.int64)
    K = int(counts.sum())
    total = (means * counts[:, None, None, None]).sum(axis=0)
    image = total / K
    # two halves of the blocks by sample count
    order = np.argsort(-counts, kind="stable")
    a, b, ca, cb = 0.0, 0.0, 0, 0
    for i in order:
        if ca <= cb:
            a = a + means[i] * counts[i]; ca += counts[i]
        else:
            b = b + means[i] * counts[i]; cb += counts[i]
    half = ou.rel_rmse(a / max(ca, 1), b / max(cb, 1)) if ca and cb else 0.0
    loo = [ou.rel_rmse(means[i], (total - means[i]) / (K - 1)) for i in range(len(counts)) if counts[i] == 1 and K > 1]
    np.savez_compressed(out, image=image.astype(np.float32), spp_total=K * S, K=K, S=S, width=w, height=h, block_means=means.astype(np.float32),
                        block_counts=counts, half_vs_half_rmse=half, half_counts=np.array([ca, cb]), loo_rmse=np.array(loo if loo else [0.0]), seconds=seconds)
    return K, half


def main():
    name, w, h, S, add = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
    threads = int(sys.argv[6]) if len(sys.argv) > 6 else -1
    out = os.path.join(ou.GOLDEN, "%s_mean_converged.npz" % name)
    means, counts, seconds = [], [], 0.0
    if os.path.exists(out):
        z = np.load(out)
        assert int(z["S"]) == S and int(z["width"]) == w and int(z["height"]) == h
        if "block_means" in z.files:
            means, counts = list(z["block_means"].astype(np.float64)), list(z["block_counts"])
        else:   # first-generation fixture: one block of K renders
            means, counts = [z["image"].astype(np.float64)], [int(z["K"])]
        seconds = float(z["seconds"])
    t0 = time.time()
    for _ in range(add):
        k = int(np.sum(counts)) if counts else 0
        y = CASES[name](w, h, S)
        y.render_params["adv_base_sampling_offset"] = ("int", k * S)
        film, info = ou.run_reference_render(y, threads=threads, flavour="ieee", timeout=24 * 3600)
        means.append(film_to_image(film)[..., :3].astype(np.float64)); counts.append(1)
        K, half = save(out, means, counts, S, w, h, seconds + time.time() - t0)
        print("render %d done (%.0f s this run), K = %d, half-vs-half rel. RMSE %.3e" % (k, time.time() - t0, K, half), flush=True)


if __name__ == "__main__":
    main()
