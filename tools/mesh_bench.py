"""Config-3 style measurement: N-triangle procedural mesh, path depth 8 (development + profiles/ evidence).
Prints one JSON line: BVH build time, incoherent-ray trace rate (rays resident in HBM), wavefront render rate and stage times."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from libyafaray_b200 import scenes, Context, capi

n_tris = int(float(sys.argv[1])) if len(sys.argv) > 1 else 10_000_000
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 2
flags = capi.FLAG_TIME_STAGES | (capi.FLAG_COUNT_TRAVERSAL if "count" in sys.argv else 0)
t0 = time.time(); y = scenes.procedural_mesh(n_tris, 1920, 1080, 16); desc, keep = y.scene_desc(); t_host = time.time() - t0
ctx = Context(0, flags)
for a in sys.argv:
    if a.startswith("leaf="):
        ctx.set_option("leaf_size", int(a[5:]))
    if a.startswith("opt:"):
        ctx.set_option(a[4:].split("=")[0], int(a.split("=")[1]))
t0 = time.time(); ctx.upload(desc, keep); t_up = time.time() - t0
st = ctx.stats()
out = {"argv": sys.argv[1:], "n_tris": int(desc.geometry.n_tris), "host_scene_s": t_host, "upload_s": t_up, "bvh_build_ms": st["ms_build"], "bvh_nodes": st["bvh_nodes"],
       "bvh_mb": st["bvh_bytes"] / 1e6, "tri_mb": st["tri_bytes"] / 1e6}
# incoherent rays: origins on a plane above the terrain, cosine-ish random directions downwards + random points/dirs inside the bound
n = 1 << 24
g = torch.Generator(device="cuda").manual_seed(1)
o = torch.rand((n, 3), device="cuda", generator=g) * torch.tensor([20.0, 20.0, 3.0], device="cuda") + torch.tensor([-10.0, -10.0, 0.0], device="cuda")
d = torch.randn((n, 3), device="cuda", generator=g); d = d / d.norm(dim=1, keepdim=True)
rays = torch.empty((n, 8), device="cuda"); rays[:, 0:3] = o; rays[:, 3] = 5e-5; rays[:, 4:7] = d; rays[:, 7] = -1.0
hits = torch.empty((n, 4), device="cuda")
stream = torch.cuda.ExternalStream(ctx.stream())
torch.cuda.synchronize()
for mode, key in ((capi.TRACE_CLOSEST, "incoherent_closest"), (capi.TRACE_SHADOW, "incoherent_shadow")):
    ctx.trace_device(rays.data_ptr(), n, hits.data_ptr(), mode); ctx.sync()
    ctx.reset_stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    reps = 5
    for _ in range(reps):
        ctx.trace_device(rays.data_ptr(), n, hits.data_ptr(), mode)
    e1.record(stream); ctx.sync(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    s2 = ctx.stats()
    out[key] = {"rays": n, "ms": ms, "mrays_s": n / ms / 1e3, "hit_frac": float((hits[:, 3].view(torch.int32) >= 0).float().mean()) if mode == 0 else float((hits[:, 0] > 0).float().mean()),
                "nodes_per_ray": (s2["nodes_visited"] + s2["nodes_visited_shadow"]) / (reps * n) if s2["nodes_visited"] + s2["nodes_visited_shadow"] else None,
                "tris_per_ray": (s2["tris_tested"] + s2["tris_tested_shadow"]) / (reps * n) if s2["tris_tested"] + s2["tris_tested_shadow"] else None}
del rays, hits, o, d
# wavefront render
ctx.film_clear(); ctx.render_pass(16, 0, sample_begin=0, sample_end=spp); ctx.sync(); ctx.reset_stats()   # warm-up: same wave size as the timed call
t0 = time.time(); ctx.render_pass(16, 0, sample_begin=spp, sample_end=2 * spp); ctx.sync(); t = time.time() - t0
s = ctx.stats()
rays_n = s["rays_camera"] + s["rays_extension"] + s["rays_shadow"] + s["rays_shadow_transp"]
out["render"] = {"spp": spp, "seconds": t, "mrays_s": rays_n / t / 1e6, "mrays_s_device": rays_n / max(1e-9, s["ms_total"]) / 1e3, "samples_s": s["camera_samples"] / t, "rays_per_sample": rays_n / max(1, s["camera_samples"]),
                 "ms_trace": s["ms_trace"], "ms_shadow": s["ms_shadow"], "ms_shade": s["ms_shade"], "ms_total": s["ms_total"], "waves": s["waves"], "retries": s["wave_retries"],
                 "rays_ext": s["rays_extension"], "rays_shadow": s["rays_shadow"],
                 "nodes_per_ray": (s["nodes_visited"] + s["nodes_visited_shadow"]) / max(1, rays_n) if s["nodes_visited"] else None,
                 "tris_per_ray": (s["tris_tested"] + s["tris_tested_shadow"]) / max(1, rays_n) if s["tris_tested"] else None,
                 "closest_nodes_per_ray": s["nodes_visited"] / max(1, s["rays_camera"] + s["rays_extension"]) if s["nodes_visited"] else None,
                 "closest_tris_per_ray": s["tris_tested"] / max(1, s["rays_camera"] + s["rays_extension"]) if s["nodes_visited"] else None,
                 "shadow_nodes_per_ray": s["nodes_visited_shadow"] / max(1, s["rays_shadow"]) if s["nodes_visited_shadow"] else None,
                 "bvh_width": s["bvh_width"], "bvh_node_bytes": s["bvh_node_bytes"], "bvh_levels": s["bvh_levels"]}
img = ctx.read_film()
out["render"]["finite"] = bool(np.isfinite(img).all()); out["render"]["mean"] = float((img[..., :3].sum(axis=(0, 1)) / img[..., 4].sum()).mean())
print(json.dumps(out))
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "mesh_%d.npz" % n_tris), film=img[::4, ::4])
