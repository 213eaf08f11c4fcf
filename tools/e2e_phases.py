"""Wall-clock split of one end-to-end bench step (upload+BVH build, film clear, render, film D2H); development aid.
usage: python tools/e2e_phases.py [NAME=INT option ...]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402
from libyafaray_b200 import capi  # noqa: E402
from libyafaray_b200.render import Context  # noqa: E402


def main():
    torch.cuda.set_device(0)
    y = bench.build_scene(1920, 1080, bench.PASS_SPP)
    desc, keep = y.scene_desc()
    ctx = Context(0)
    for kv in sys.argv[1:]:
        ctx.set_option(kv.split("=")[0], int(kv.split("=")[1]))
    host_film = torch.empty((1080, 1920, 5), dtype=torch.float32, pin_memory=True)
    tot = {}
    for i in range(10):
        t = [time.perf_counter()]
        ctx.upload(desc, keep); t.append(time.perf_counter())
        ctx.film_clear(); ctx.sync(); t.append(time.perf_counter())
        ctx.render_pass(bench.PASS_SPP, 0, False, 0, sample_begin=(i % 64) * 4, sample_end=(i % 64) * 4 + 4); ctx.sync(); t.append(time.perf_counter())
        capi.lib.yafgpu_read_film(ctx.h, host_film.data_ptr()); t.append(time.perf_counter())
        if i >= 2:
            for k, name in enumerate(("upload", "clear", "render", "read_film")):
                tot[name] = tot.get(name, 0.0) + (t[k + 1] - t[k]) / 8 * 1e3
    print({k: round(v, 3) for k, v in tot.items()}, "ms per step")


if __name__ == "__main__":
    main()
