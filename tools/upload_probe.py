import sys, time
sys.path.insert(0, '/root/repo')
from libyafaray_b200 import scenes, Context
y = scenes.procedural_mesh(10_000_000, 1920, 1080, 16)
desc, keep = y.scene_desc()
ctx = Context(0)
for i in range(3):
    t = time.time(); ctx.upload(desc, keep); print("upload %d: %.1f ms" % (i, (time.time() - t) * 1e3), file=sys.stderr)
