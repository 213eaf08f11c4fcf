"""Phases of yafgpu_upload_scene on the bench scene (YAFGPU_UPLOAD_TIMING=1 prints them), pageable and page-locked host arrays."""
import os
import sys
import time

sys.path.insert(0, "/root/repo")
os.environ["YAFGPU_UPLOAD_TIMING"] = "1"
from libyafaray_b200 import scenes, Context, pin_host_arrays  # noqa: E402

y = scenes.procedural_mesh(10_000_000, 1920, 1080, 16, threads=-1)
desc, keep = y.scene_desc()
ctx = Context(0)
for label in ("pageable", "pinned"):
    if label == "pinned":
        pin_host_arrays(keep)
    for i in range(3):
        t = time.time(); ctx.upload(desc, keep); ctx.sync()
        print("%s upload %d: %.1f ms" % (label, i, (time.time() - t) * 1e3), file=sys.stderr, flush=True)
