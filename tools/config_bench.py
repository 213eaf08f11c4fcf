"""BASELINE.json configs[2..4] on N GPUs of one box (development + profiles/ evidence; bench.py stays the contract bench on configs[1]).

  python tools/config_bench.py c3            # 10^7-triangle mesh, glossy/coated/rough glass, depth 8, 1 GPU
  torchrun --nproc-per-node 4 tools/config_bench.py c4   # 10^6 triangles + HDR IBL (ibl_samples 16) + 3 mesh lights + blend materials, 1920x1080
  torchrun --nproc-per-node 8 tools/config_bench.py c5   # 3840x2160, 2*10^7 triangles, tile-sharded, NCCL film reduce

Every rank holds the whole scene, renders `spp` samples per pixel of its own interleaved 32x32 tiles per step (weak scaling: the frame advances N x
faster), and the step ends with the NCCL sum of the w*h*5 film on rank 0.  Timing: CUDA events on the context's stream around K steps after W warm-up
steps, barrier + synchronize on both sides, max over ranks.  Prints one JSON line on rank 0."""
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from libyafaray_b200 import scenes, Context, capi  # noqa: E402
from libyafaray_b200.interface import write_radiance_hdr  # noqa: E402

CONFIGS = {
    "c3": dict(n_tris=10_000_000, width=1920, height=1080, pass_spp=16, spp=2),
    "c4": dict(n_tris=1_000_000, width=1920, height=1080, pass_spp=1024, spp=2, ibl=True, mesh_lights=3, blend=True),
    "c5": dict(n_tris=20_000_000, width=3840, height=2160, pass_spp=4096, spp=1, seed=23456),
}


def synthetic_env(path, w=256, h=128):
    """Seeded equirectangular HDR: sky gradient, a bright sun disc and a few coloured lobes (no network for real HDRIs)."""
    v, u = np.mgrid[0:h, 0:w].astype(np.float32)
    theta, phi = (v + 0.5) / h * np.pi, (u + 0.5) / w * 2 * np.pi
    d = np.stack([np.sin(theta) * np.cos(phi), np.sin(theta) * np.sin(phi), np.cos(theta)], axis=-1)
    img = np.empty((h, w, 3), np.float32)
    up = np.clip(d[..., 2], 0, 1)
    img[...] = (0.25 + 0.6 * up)[..., None] * np.array([0.5, 0.7, 1.0], np.float32) + (1 - up)[..., None] * 0.08
    for c, col, sharp, gain in (((0.4, -0.5, 0.75), (1.0, 0.9, 0.7), 400.0, 60.0), ((-0.7, 0.2, 0.3), (0.9, 0.3, 0.2), 20.0, 2.0), ((0.1, 0.9, 0.2), (0.2, 0.8, 0.4), 12.0, 1.5)):
        c = np.array(c, np.float32); c /= np.linalg.norm(c)
        img += (gain * np.exp(sharp * (d @ c - 1.0)))[..., None] * np.array(col, np.float32)
    write_radiance_hdr(path, img)
    return path


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c3"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    warmup = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    cfg = dict(CONFIGS[name])
    for a in sys.argv[4:]:
        k, v = a.split("="); cfg[k] = int(float(v))
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    td = tempfile.mkdtemp()
    env = synthetic_env(os.path.join(td, "env_%d.hdr" % rank)) if cfg.get("ibl") else None
    y = scenes.procedural_mesh(cfg["n_tris"], cfg["width"], cfg["height"], cfg["pass_spp"], seed=cfg.get("seed", 12345), ibl_hdr=env,
                               mesh_lights=cfg.get("mesh_lights", 0), blend=cfg.get("blend", False))
    desc, keep = y.scene_desc()
    ctx = Context(local, capi.FLAG_TIME_STAGES)
    ctx.upload(desc, keep)
    if "leaf_size" in cfg:   # A/B knob: python tools/config_bench.py c3 4 2 leaf_size=4 (before the upload below? no: re-upload)
        ctx.set_option("leaf_size", cfg["leaf_size"]); ctx.upload(desc, keep)
    st0 = ctx.stats()
    S = cfg["spp"] * world
    n_chunks = max(1, cfg["pass_spp"] // S)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=local)
    fptr, fcount = ctx.film_device_ptr()

    class _Film:
        __cuda_array_interface__ = {"shape": (fcount,), "typestr": "<f4", "data": (fptr, False), "version": 3, "strides": None}
    film = torch.as_tensor(_Film(), device="cuda:%d" % local)
    red = torch.empty_like(film) if world > 1 else None

    def step(i):
        b = (i % n_chunks) * S
        ctx.render_pass(cfg["pass_spp"], 0, False, 0, sample_begin=b, sample_end=min(cfg["pass_spp"], b + S), part_index=rank, part_count=world)
        if world > 1:
            red.copy_(film)
            dist.reduce(red, 0, op=dist.ReduceOp.SUM)

    ctx.film_clear()
    for i in range(warmup):
        step(i)
    ctx.reset_stats()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(steps):
        step(warmup + i)
    if world > 1:
        torch.cuda.current_stream().synchronize()
    e1.record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    st = ctx.stats()
    rays = float(st["rays_camera"] + st["rays_extension"] + st["rays_shadow"] + st["rays_shadow_transp"] + st["rays_light_probe"])
    vals = torch.tensor([ms, rays, float(st["camera_samples"]), st["ms_trace"], st["ms_shadow"], st["ms_shade"]], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = vals.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = vals.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        ms, rays, samples = float(mx[0]), float(sm[1]), float(sm[2])
    else:
        samples = float(vals[2])
    img = ctx.read_film() if world == 1 else red.cpu().numpy().reshape(cfg["height"], cfg["width"], 5)
    if rank == 0:
        wsum = float(img[..., 4].sum())
        print(json.dumps({"config": name, "n_gpus": world, "n_tris": int(desc.geometry.n_tris), "width": cfg["width"], "height": cfg["height"], "spp_per_step_per_gpu": cfg["spp"],
                          "steps": steps, "warmup": warmup, "ms_per_step": ms / steps, "mrays_s": rays / ms / 1e3, "samples_s": samples / ms * 1e3,
                          "rays_per_sample": rays / max(1.0, samples), "bvh_build_ms": st0["ms_build"], "stage_ms_rank0": {"trace": st["ms_trace"], "shadow": st["ms_shadow"], "shade": st["ms_shade"]},
                          "film_reduce_mb": fcount * 4 / 1e6 if world > 1 else 0.0, "finite": bool(np.isfinite(img).all()),
                          "mean_rgb": [float(img[..., k].sum() / max(wsum, 1e-9)) for k in range(3)], "lights": int(desc.n_lights), "materials": int(desc.n_materials)}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
