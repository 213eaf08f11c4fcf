"""Exploratory GPU-vs-oracle comparison (development tool; prints one JSON line per check, saves films to gpurun_out/)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libyafaray_b200 import scenes, Context, render_scene, film_to_image, capi  # noqa: E402
from tests import oracle_util as ou  # noqa: E402

OUT = os.path.join(ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)


def random_rays(n, lo, hi, seed=1):
    rng = np.random.default_rng(seed)
    o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = np.zeros((n, 8), np.float32)
    r[:, 0:3], r[:, 3], r[:, 4:7], r[:, 7] = o, 5e-5, d, -1.0
    return r


def check_trace(name, y, ctx, n=200000):
    desc, keep = y.scene_desc()
    ctx.upload(desc, keep)
    b = ctx.scene_bound()
    ext = b[3:] - b[:3]
    rays = random_rays(n, b[:3] - 0.2 * ext, b[3:] + 0.2 * ext)
    for mode in (0, 1):
        r = rays.copy()
        if mode == 1:
            r[:, 3] = 5e-4
            r[:, 7] = np.random.default_rng(5).uniform(0.1, 1.5 * float(np.linalg.norm(ext)), n).astype(np.float32)
        t0 = time.time(); g = ctx.trace(r, mode); tg = time.time() - t0
        t0 = time.time(); h = ou.reference_hits(y, r, mode); tr = time.time() - t0
        if h is None:
            print(json.dumps({"check": "trace", "scene": name, "mode": mode, "oracle": "missing"})); continue
        if mode == 0:
            same = g["tri"] == h["tri"]
            bits = same & (g["t"].view(np.uint32) == h["t"].view(np.uint32)) & (g["u"].view(np.uint32) == h["u"].view(np.uint32))
            bad = np.flatnonzero(~same)
            tie = 0
            for i in bad[:1000]:
                if g["tri"][i] >= 0 and h["tri"][i] >= 0 and abs(g["t"][i] - h["t"][i]) <= 1e-6 * max(1.0, abs(h["t"][i])):
                    tie += 1
            print(json.dumps({"check": "trace_closest", "scene": name, "n": n, "id_mismatch": int(len(bad)), "ties": tie,
                              "tuv_bit_exact": int(bits.sum()), "hits": int((h["tri"] >= 0).sum()), "gpu_s": tg, "oracle_s": tr}))
            if len(bad):
                i = bad[0]
                print("  first mismatch", i, r[i], g[i], h[i])
        else:
            sg, sh = g["tri"] >= 0, h["tri"] >= 0
            print(json.dumps({"check": "trace_shadow", "scene": name, "n": n, "mismatch": int((sg != sh).sum()), "shadowed": int(sh.sum())}))


def check_render(name, y, ctx, ref_y=None, save=True):
    t0 = time.time(); film = render_scene(y, ctx=ctx); tg = time.time() - t0
    st = ctx.stats()
    ref, info = ou.run_reference_render(ref_y or y)
    img = film_to_image(film)
    rec = {"check": "render", "scene": name, "gpu_s": round(tg, 3), "mean_gpu": [float(x) for x in img[..., :3].mean(axis=(0, 1))],
           "finite": bool(np.isfinite(img).all()), "waves": st["waves"], "retries": st["wave_retries"],
           "rays": st["rays_camera"] + st["rays_extension"] + st["rays_shadow"] + st["rays_shadow_transp"], "ms_total": st["ms_total"]}
    if ref is not None:
        rimg = film_to_image(ref)
        d = img[..., :3] - rimg[..., :3]
        rec.update({"mean_ref": [float(x) for x in rimg[..., :3].mean(axis=(0, 1))], "rmse": float(np.sqrt(np.mean(d * d))),
                    "rel_rmse": ou.rel_rmse(img, rimg), "max_abs": float(np.abs(d).max()), "weight_max_diff": float(np.abs(film[..., 4] - ref[..., 4]).max()),
                    "ref_s": info.get("rendert_s")})
        if save:
            np.savez_compressed(os.path.join(OUT, name + ".npz"), gpu=film, ref=ref)
    ctx.reset_stats()
    print(json.dumps(rec)); sys.stdout.flush()
    return rec


def main():
    ctx = Context(0)
    which = sys.argv[1:] or ["trace", "simple", "cornell_dl", "cornell_pt"]
    if "trace" in which:
        check_trace("simple", scenes.simple_point_scene(), ctx)
        check_trace("cornell", scenes.cornell_box(width=64, height=36, spp=1), ctx)
    if "simple" in which:
        check_render("simple_dl_gauss", scenes.simple_point_scene(160, 90, 1), ctx)
        check_render("simple_dl_box4", scenes.simple_point_scene(160, 90, 4, filter_type="box"), ctx)
        check_render("simple_pt", scenes.simple_point_scene(160, 90, 8, integrator="pathtracing", filter_type="box"), ctx)
    if "cornell_dl" in which:
        check_render("cornell_dl", scenes.cornell_box(160, 90, 4, integrator="directlighting"), ctx)
        check_render("cornell_dl_noglass", scenes.cornell_box(160, 90, 4, integrator="directlighting", glass=False), ctx)
    if "cornell_pt" in which:
        check_render("cornell_pt_noglass_64", scenes.cornell_box(160, 90, 64, glass=False, bounces=4), ctx)
        check_render("cornell_pt_64", scenes.cornell_box(160, 90, 64, bounces=4, raydepth=4), ctx)
        check_render("cornell_pt_ps4", scenes.cornell_box(160, 90, 16, path_samples=4, bounces=8, raydepth=8), ctx)
    ctx.close()


if __name__ == "__main__":
    main()
