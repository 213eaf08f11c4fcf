"""Device-counted rays per camera sample of the bench workload, for the full frame and for the centred half-size window the CPU
arm renders (bench.py crop_window): writes profiles/rays_per_sample.json.  The reference has no ray counter, so its Mrays/s is
samples/s x these factors (labelled derived in every line that uses them).

    python tools/rays_per_sample.py [--tris N]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from libyafaray_b200 import Context  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=bench.N_TRIS)
    ap.add_argument("--width", type=int, default=bench.WIDTH)
    ap.add_argument("--height", type=int, default=bench.HEIGHT)
    args = ap.parse_args()
    out = {"workload": bench.config_dict(args, 1, bench.FRAME_SPP)["workload"]}
    ctx = Context(0)
    for key, crop, spp in (("frame", None, bench.FRAME_SPP), ("crop", bench.crop_window(args.width, args.height), 1), ("frame_1spp", None, 1)):
        y = bench.build_scene(args, spp, crop=crop)
        desc, keep = y.scene_desc()
        ctx.upload(desc, keep)
        ctx.film_clear()
        ctx.reset_stats()
        ctx.render_pass(spp, 0)
        st = ctx.stats()
        out[key] = bench.total_rays(st) / st["camera_samples"]
        out[key + "_camera_samples"] = st["camera_samples"]
    ctx.close()
    with open(bench.RAYS_FILE, "w") as fh:
        json.dump(out, fh, indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
