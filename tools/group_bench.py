"""The bench workload through the IN-PROCESS multi-GPU path (yafgpu_group: one process, N GPUs, film sum inside libyafgpu.so) — the
drop-in's path, next to `bench.py --gpus N` (one process per GPU, NCCL).  Weak scaling like bench.py: a step adds 16 x N spp.
usage: python tools/group_bench.py [n_gpus] [steps] [warmup]   -> one JSON line"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from libyafaray_b200 import Group  # noqa: E402


class A:
    tris, width, height = bench.N_TRIS, bench.WIDTH, bench.HEIGHT


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    warmup = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    S = bench.FRAME_SPP * n
    y = bench.build_scene(A, S)
    desc, keep = y.scene_desc()
    share = "share" in sys.argv   # n members on GPU 0 (the partition + film-sum path on a one-GPU box)
    from libyafaray_b200 import capi
    g = Group([0] * n, capi.FLAG_SHARE_DEVICE) if share else Group(list(range(n)))
    t0 = time.perf_counter(); g.upload(desc, keep); t_up = time.perf_counter() - t0
    g.film_clear()
    for i in range(warmup):
        g.render_pass(S, i * S)
    g.sync(); g.reset_stats()
    t0 = time.perf_counter()
    for i in range(steps):
        g.render_pass(S, (warmup + i) * S)
    film = g.read_film()     # the summed film of the last step; the earlier sums overlap the following pass
    t = time.perf_counter() - t0
    st = g.stats()
    rays = bench.total_rays(st)
    print(json.dumps({"path": "yafgpu_group (in-process)", "n_gpus": n, "steps": steps, "warmup": warmup, "ms_per_step": t / steps * 1e3, "value_mrays_s": rays / t / 1e6,
                      "samples_per_s": st["camera_samples"] / t, "upload_s": t_up, "film_reduces": st["film_reduces"], "ms_last_film_reduce": st["ms_film_reduce"],
                      "timing": "host wall clock around `steps` group passes + the final film read-back", "finite": bool(np.isfinite(film).all()), "non_finite_pixels": int((~np.isfinite(film).all(axis=2)).sum()), "first_bad": np.argwhere(~np.isfinite(film).all(axis=2))[:4].tolist(),
                      "workload": bench.config_dict(A, n, bench.FRAME_SPP)["workload"]}))
    if "members" in sys.argv:   # diagnosis: each member's own film next to the sum
        g.sync()
        tot = np.zeros_like(film, np.float64)
        for k in range(g.size()):
            m = np.empty_like(film)
            assert capi.lib.yafgpu_read_film(capi.lib.yafgpu_group_ctx(g.h, k), m.ctypes.data) == 0
            bad = ~np.isfinite(m).all(axis=2)
            print(f"member {k}: non-finite pixels {int(bad.sum())} first {np.argwhere(bad)[:3].tolist()} weight sum {float(m[..., 4][~bad].sum()):.1f}", file=sys.stderr)
            tot += np.nan_to_num(m.astype(np.float64), nan=0.0, posinf=0.0, neginf=0.0)
        ok = np.isfinite(film).all(axis=2)
        print("sum of member films vs reduced film (finite pixels): max |d| =", float(np.abs(tot[ok] - film[ok]).max()), file=sys.stderr)
    g.close()


if __name__ == "__main__":
    main()
