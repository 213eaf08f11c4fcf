"""Benchmark of the path-tracing hot path (contract: python bench.py --gpus N --steps K --warmup W [--impl reference]).

Workload = BASELINE.json configs[2], the largest single-GPU configuration and the one `north_star`'s >= 1 Grays/s target is
quoted on: the synthetic 10^7-triangle procedural mesh (glossy / coated_glossy / rough_glass / shinydiffuse patches, sun + area
light), PathIntegrator with path depth 8 (bounces 8, raydepth 8, russian_roulette_min_bounces 2, path_samples 1), 1920x1080,
16 samples per pixel per frame.  A "step" is ONE FRAME of that configuration per GPU: `--spp-per-step`
(16) samples per pixel of this rank's interleaved 32x32 tiles, scene replicated; on N GPUs a step therefore adds 16 x N spp to
the frame (per-GPU work fixed -> weak scaling) and ends with the sum of the film accumulators on rank 0 (the film is a linear
accumulator).  Successive steps use successive sampling offsets, like successive AA passes.

  value          whole-job Mrays/s (all ray kinds, counted on the device) with the scene resident in HBM
  samples_per_s  camera samples/s (the reference has no ray counter; its Mrays/s is samples/s x the device-counted rays per
                 camera sample of this very workload, labelled derived, `rays_source` says where the factor comes from)
  e2e            the same metric through the public call with HOST buffers, every step: yafgpu_upload_scene (0.5 GB of geometry
                 H2D + BVH build), film clear, render of the frame, film read-back (41.5 MB D2H)
  roofline       the kernel class with the largest share of the step (closest-hit traversal on this scene): algorithmic bytes
                 (per ray 48 B ray+hit record + node bytes x nodes visited + 48 B x triangles tested, counted by a
                 COUNT_TRAVERSAL context on the same steps) / its duration from CUDA events on the launching stream
  cpu_baseline   the reference's own CPU code (oracle/_ref/yaf_ref_fast, reference Release flags, all host threads) on a
                 bounded sample of the same scene: the full frame at 1 spp per step (--ref-crop: the centred half-size window)

`--impl reference` runs only that CPU arm (scene description -> XML through the streaming exporter -> the reference's unchanged XML
loader -> kd-tree build -> Scene::render), in a process that never maps libyafgpu.so.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_TRIS, WIDTH, HEIGHT, FRAME_SPP = 10_000_000, 1920, 1080, 16
RAYS_FILE = os.path.join(ROOT, "profiles", "rays_per_sample.json")
REF_CACHE = os.path.join(tempfile.gettempdir(), "yafgpu_bench_reference_arm.json")


def crop_window(width, height):
    """The bounded CPU sample: centred window of half the frame in each dimension (a quarter of the pixels)."""
    return width // 4, height // 4, width // 2, height // 2


def build_scene(args, spp, crop=None):
    from libyafaray_b200 import scenes
    return scenes.procedural_mesh(args.tris, args.width, args.height, spp, seed=12345, bounces=8, raydepth=8, path_samples=1, rr_min_bounces=2,
                                  integrator="pathtracing", crop=crop, threads=-1)


def config_dict(args, world, spp_per_step):
    return {"workload": "procedural_mesh_%dM_triangles_pathtracing_depth8_%dx%d_%dspp_frame" % (round(args.tris / 1e6), args.width, args.height, FRAME_SPP),
            "baseline_config": "configs[2]: synthetic 10M-triangle procedural mesh, Glossy/CoatedGlossy/RoughGlass, path depth 8, 1 GPU",
            "triangles": args.tris, "integrator": "pathtracing", "path_samples": 1, "bounces": 8, "raydepth": 8, "rr_min_bounces": 2,
            "lights": "sunlight + arealight, samples=1", "materials": "glossy x2, coated_glossy x2, rough_glass, shinydiffusemat, light_mat",
            "spp_per_step_per_gpu": spp_per_step, "camera_samples_per_step": args.width * args.height * spp_per_step * world,
            "partition": "8x8-pixel blocks interleaved x%d (shifted rows), scene replicated" % world,
            "cache": "working set (1.1 GB of BVH + triangles, > 1 GB of wavefront queues per wave) exceeds the 126 MB L2; no explicit flush"}


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.rows = index, False, []

    def run(self):
        # NVML in-process (a sample every 10 ms, so that sub-second timed regions are covered); nvidia-smi as the fallback
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            R = pynvml
            bits = [("hw_slowdown", R.nvmlClocksThrottleReasonHwSlowdown), ("hw_thermal_slowdown", R.nvmlClocksThrottleReasonHwThermalSlowdown),
                    ("sw_thermal_slowdown", R.nvmlClocksThrottleReasonSwThermalSlowdown), ("sw_power_cap", R.nvmlClocksThrottleReasonSwPowerCap)]
            while not self.stop_flag:
                sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                rs = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.rows.append([str(sm), str(mx)] + ["Active" if rs & b else "Not Active" for _, b in bits])
                time.sleep(0.01)
            return
        except Exception:
            pass
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, timeout=5).stdout.decode().strip()
                if out:
                    self.rows.append([x.strip() for x in out.splitlines()[0].split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.rows[0][1]) if self.rows[0][1].replace(".", "").isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def rays_per_sample(key):
    """Device-counted rays per camera sample of this workload, committed by a builder run of tools/rays_per_sample.py (the CPU
    reference has no ray counter).  key: "frame" (16 spp), "frame_1spp" (what the CPU arm renders) or "crop" (--ref-crop)."""
    try:
        with open(RAYS_FILE) as fh:
            d = json.load(fh)
        return float(d[key]), "profiles/rays_per_sample.json[%s], device-counted by bench.py on %s" % (key, d.get("workload", "?"))
    except Exception:
        return 16.9, "constant 16.9 (profiles/rays_per_sample.json missing)"


def run_cpu_reference(args, warmup, steps, timeout=3000):
    """oracle/_ref/yaf_ref_fast bench: XML load + kd-tree build once, then warmup + steps renders of the bounded sample."""
    exe = os.path.join(ROOT, "oracle", "_ref", "yaf_ref_fast")
    if not os.path.exists(exe):
        return None
    crop = crop_window(args.width, args.height) if args.ref_crop else (0, 0, args.width, args.height)
    t0 = time.time()
    y = build_scene(args, args.ref_spp, crop=crop if args.ref_crop else None)
    with tempfile.TemporaryDirectory() as td:
        xml = os.path.join(td, "scene.xml")
        y.to_xml(xml)
        t_xml = time.time() - t0
        r = subprocess.run([exe, "bench", xml, "-1", str(warmup), str(steps), str(args.ref_spp)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, timeout=timeout)
    for line in reversed(r.stdout.decode(errors="replace").splitlines()):
        if line.startswith("{"):
            info = json.loads(line)
            info["scene_and_xml_export_s"] = t_xml
            info["crop"] = crop
            return info
    return None


def cpu_baseline_dict(args, info, rps, rps_src):
    x0, y0, w, h = info["crop"]
    t = sum(info["rendert_s"]) / len(info["rendert_s"])
    sps = w * h * info["spp"] / t
    return {"value": sps * rps / 1e6, "unit": "Mrays/s", "cores": int(info["threads"]), "kind": "reference",
            "sample": ("the full %dx%d frame" % (w, h) if (w, h) == (args.width, args.height) else "centred %dx%d window (origin %d,%d) of the %dx%d frame" % (w, h, x0, y0, args.width, args.height))
                      + " at %d spp per step (the GPU step is %d spp), same scene file, sampler and integrator settings; extrapolated: Mrays/s = CPU samples/s x device-counted rays per camera sample" % (info["spp"], FRAME_SPP),
            "samples_per_s": sps, "rays_per_camera_sample": rps, "rays_source": rps_src, "rendert_s_per_step": t,
            "kdtree_build_s": info["scene_update_s"], "xml_parse_s": info["xml_parse_s"], "scene_and_xml_export_s": info["scene_and_xml_export_s"],
            "note": "render time = the reference's own rendert timer (integrator_tiled.cc:168-169,253-255); kd-tree build (Scene::update, kdtree_triangle.cc:76-157) "
                    "and XML parse are reported separately and NOT included; global film mutex (imagefilm.cc:960)"}


def reference_arm(args, rank, world):
    """--impl reference: the reference's own CPU implementation on the host cores (rank 0 only).  Never maps libyafgpu.so."""
    if rank != 0:
        return
    os.environ["LIBYAFARAY_B200_HOST_ONLY"] = "1"   # scene description + XML export only (libyafaray_b200/capi.py)
    info = run_cpu_reference(args, args.warmup, args.steps)
    if info is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/yaf_ref_fast is not built"}))
        return
    rps, rps_src = rays_per_sample("crop" if args.ref_crop else "frame_1spp")
    cb = cpu_baseline_dict(args, info, rps, rps_src)
    cfg = config_dict(args, world, args.ref_spp)
    x0, y0, w, h = info["crop"]
    cfg["spp_per_step_per_gpu"] = None
    cfg["camera_samples_per_step"] = w * h * info["spp"]
    cfg["reference_step"] = cb["sample"]
    line = {"impl": "reference", "metric": "Mrays/s", "value": cb["value"], "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": cb["rendert_s_per_step"] * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "samples_per_s": cb["samples_per_s"], "rays_per_camera_sample": rps, "rays_derived": True, "rays_source": rps_src, "config": cfg,
            "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "native_so": "libyafgpu" in open("/proc/self/maps").read()}
    try:
        with open(REF_CACHE, "w") as fh:
            json.dump({"tris": args.tris, "width": args.width, "height": args.height, "ref_spp": args.ref_spp, "ref_crop": bool(args.ref_crop), "info": info}, fh)
    except OSError:
        pass
    print(json.dumps(line))


def total_rays(st):
    return st["rays_camera"] + st["rays_extension"] + st["rays_shadow"] + st["rays_shadow_transp"] + st["rays_light_probe"]


def gpu_arm(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    from libyafaray_b200 import Context, capi

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    S = args.spp_per_step * world            # samples per pixel one step adds to the frame
    y = build_scene(args, S)
    desc, keep = y.scene_desc()
    ctx = Context(local_rank)
    if args.wave_size:
        ctx.set_option("wave_size", args.wave_size)
    for kv in args.option:
        ctx.set_option(kv.split("=")[0], int(kv.split("=")[1]))
    ctx.upload(desc, keep)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=local_rank)

    def wrap_film(c):
        fptr, fcount = c.film_device_ptr()

        class _Film:
            __cuda_array_interface__ = {"shape": (fcount,), "typestr": "<f4", "data": (fptr, False), "version": 3, "strides": None}
        return torch.as_tensor(_Film(), device="cuda:%d" % local_rank)
    films = {}
    red = torch.empty_like(wrap_film(ctx)) if world > 1 else None

    def step(i, c=ctx):
        # one frame: this rank's tiles at S samples per pixel, sampling offset advancing like successive AA passes
        c.render_pass(S, i * S, False, 0, part_index=rank, part_count=world)
        if world > 1:
            if id(c) not in films:
                films[id(c)] = wrap_film(c)
            with torch.cuda.stream(stream):   # ordered after the pass on the context's own stream
                red.copy_(films[id(c)])
                dist.reduce(red, 0, op=dist.ReduceOp.SUM)

    ctx.film_clear()
    for i in range(args.warmup):
        step(i)
    ctx.reset_stats()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(args.steps):
        step(args.warmup + i)
    e1.record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    st = ctx.stats()
    rays, samples = total_rays(st), st["camera_samples"]
    launches = st["kernel_launches"]
    if world > 1:
        t = torch.tensor([ms, float(rays), float(samples), float(launches)], dtype=torch.float64, device="cuda")
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms, rays, samples, launches = float(tmax[0]), float(tsum[1]), float(tsum[2]), int(tsum[3])
    sec = ms / 1e3
    line = {"metric": "Mrays/s", "value": rays / sec / 1e6, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "samples_per_s": samples / sec, "rays_per_camera_sample": rays / max(1.0, samples), "config": config_dict(args, world, args.spp_per_step),
            "gpu_launches": int(launches), "clocks": sampler.summary(), "waves_per_step": st["waves"] / max(1, args.steps), "wave_retries": st["wave_retries"]}

    # ---- end-to-end through the public call with host buffers (every step: upload + BVH build, clear, render, film D2H) ----
    host_film = torch.empty((args.height, args.width, 5), dtype=torch.float32, pin_memory=True)  # pinned: the D2H of the film runs at PCIe speed
    from libyafaray_b200 import pin_host_arrays
    pinned = pin_host_arrays(keep) if not args.pageable else []   # once, outside the timed region: the step's inputs come from pinned host memory
    g = desc.geometry
    h2d = (g.n_points * 12 + g.n_normals * 12 + g.n_uv * 8 + g.n_tris * (12 * (1 + (1 if g.n_normals else 0) + (1 if g.n_uv else 0)) + 8)
           + desc.n_materials * C_sizeof(capi.Material) + desc.n_lights * C_sizeof(capi.Light))
    e2e_steps = max(2, min(args.steps, 6))
    phases = {"upload_ms": 0.0, "render_ms": 0.0, "readback_ms": 0.0}

    def e2e_step(i):
        t0 = time.perf_counter()
        ctx.upload(desc, keep)
        films.pop(id(ctx), None)  # the film accumulators were re-allocated by the upload
        ctx.film_clear()
        t1 = time.perf_counter()
        step(i)
        if world > 1:
            stream.synchronize()
        t2 = time.perf_counter()
        if world > 1:
            if rank == 0:   # the result of a multi-GPU step is the REDUCED film
                host_film.view(-1).copy_(red, non_blocking=False)
        else:
            capi.lib.yafgpu_read_film(ctx.h, host_film.data_ptr())
        t3 = time.perf_counter()
        phases["upload_ms"] += (t1 - t0) * 1e3; phases["render_ms"] += (t2 - t1) * 1e3; phases["readback_ms"] += (t3 - t2) * 1e3
    e2e_step(0)
    for k in phases:
        phases[k] = 0.0
    ctx.reset_stats()
    build0 = ctx.stats()["ms_build"]   # accumulates over uploads: report the mean of the timed ones
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        e2e_step(1 + i)
    torch.cuda.synchronize()
    t_e2e = time.perf_counter() - t0
    st2 = ctx.stats()
    r2 = float(total_rays(st2))
    if world > 1:
        t = torch.tensor([t_e2e], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); t_e2e = float(t[0])
        t = torch.tensor([r2], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.SUM); r2 = float(t[0])
    line["e2e"] = {"value": r2 / t_e2e / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(host_film.numel() * 4),
                   "ms_per_step": t_e2e / e2e_steps * 1e3, "steps": e2e_steps, "phases_ms_per_step_rank0": {k: v / e2e_steps for k, v in phases.items()},
                   "bvh_build_ms": (st2["ms_build"] - build0) / max(1, e2e_steps),
                   "host_buffers": "pageable" if args.pageable else "pinned once before the timed steps (yafgpu_host_register, %d arrays)" % len(pinned),
                   "timing": "host wall clock around yafgpu_upload_scene (H2D + BVH build) + film clear + render of the frame + film read-back, every step, max over ranks"}

    if rank == 0:
        # ---- roofline of the dominant kernel class: counting + stage-timing contexts on the same steps ----
        def side_run(flags):
            c2 = Context(local_rank, flags)
            if args.wave_size:
                c2.set_option("wave_size", args.wave_size)
            for kv in args.option:
                c2.set_option(kv.split("=")[0], int(kv.split("=")[1]))
            c2.upload(desc, keep)
            c2.film_clear()
            step_plain = lambda i: c2.render_pass(S, i * S, False, 0, part_index=rank, part_count=world)
            step_plain(0)
            c2.reset_stats()
            for i in range(min(args.steps, 3)):
                step_plain(args.warmup + i)
            r = c2.stats()
            c2.close()
            return r
        ctx.close(); ctx = None                     # HBM for the side contexts
        s2 = side_run(capi.FLAG_TIME_STAGES)        # kernel durations (CUDA events around every launch, no counting code)
        s3 = side_run(capi.FLAG_COUNT_TRAVERSAL)    # nodes visited / triangles tested on the same steps
        for k in ("nodes_visited", "tris_tested", "nodes_visited_shadow", "tris_tested_shadow"):
            s2[k] = s3[k]
        line["wave0_tasks"], line["wave0_shadow"] = s2["last_wave_tasks"], s2["last_wave_shadow"]
        peak, how = hbm_peak()
        n_trace, n_shadow = s2["rays_in_timed_trace"], s2["rays_in_timed_shadow"]
        ms_tr, ms_sh, ms_shd = s2["ms_trace"], s2["ms_shadow"], s2["ms_shade"]
        # Algorithmic bytes.  Traversal: 48 B ray+hit record, node_bytes per node visited, 48 B per
        # triangle tested, closest-hit and shadow rays counted separately.  Shading: per task 108 B queue record in (ray_o, ray_d, hit,
        # wgt, thr|dc, st, perm), 104 B per task spawned, 48 B per shadow ray queued, 136 B of triangle/vertex/normal data per hit
        # (the scene exceeds the caches).
        node_b, tri_b = float(s2["bvh_node_bytes"]), 48.0
        n_rays = max(1, n_trace + n_shadow)
        per_ray_c = (node_b * s2["nodes_visited"] + tri_b * s2["tris_tested"]) / max(1, n_trace)
        per_ray_s = (node_b * s2["nodes_visited_shadow"] + tri_b * s2["tris_tested_shadow"]) / max(1, n_shadow)
        spawned = max(0, n_trace - s2["camera_samples"])
        alg_shade = (108.0 + 136.0) * n_trace + 104.0 * spawned + 48.0 * n_shadow
        kern = {
            "k_trace": {"ms": ms_tr, "launches": s2["launches_trace"], "alg_bytes": (48.0 + per_ray_c) * n_trace, "units": n_trace, "unit": "ray",
                        "nodes_per_ray": s2["nodes_visited"] / max(1, n_trace), "tris_per_ray": s2["tris_tested"] / max(1, n_trace)},
            "k_shadow": {"ms": ms_sh, "launches": s2["launches_shadow"], "alg_bytes": (48.0 + per_ray_s) * n_shadow, "units": n_shadow, "unit": "ray",
                         "nodes_per_ray": s2["nodes_visited_shadow"] / max(1, n_shadow), "tris_per_ray": s2["tris_tested_shadow"] / max(1, n_shadow)},
            "k_shade": {"ms": ms_shd, "launches": s2["launches_shade"], "alg_bytes": alg_shade, "units": n_trace, "unit": "task"},
        }
        dom = max(kern, key=lambda k: kern[k]["ms"])
        for k, v in kern.items():
            v["achieved_gbs"] = v["alg_bytes"] / (v["ms"] / 1e3) / 1e9 if v["ms"] > 0 else 0.0
            v["frac"] = v["achieved_gbs"] / peak
            v["avg_launch_ms"] = v["ms"] / max(1, v["launches"])
            v["alg_bytes_per_unit"] = v["alg_bytes"] / max(1, v["units"])
            v["share_of_step"] = v["ms"] / max(1e-9, s2["ms_total"])
        traffic = None
        try:   # dram bytes per unit of the dominant kernel from the committed ncu capture of this workload (profiles/)
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as fh:
                t = json.load(fh).get(dom)
            if t:
                traffic = t["dram_bytes_per_unit"] * kern[dom]["units"] / max(1, kern[dom]["launches"])
        except Exception:
            pass
        d = kern[dom]
        line["roofline"] = {"bound": "hbm", "achieved": d["achieved_gbs"], "peak": peak, "unit": "GB/s", "frac": d["frac"], "traffic": traffic,
                            "peak_source": how, "kernel": dom, "algorithmic_bytes_per_launch": d["alg_bytes"] / max(1, d["launches"]),
                            "avg_launch_ms": d["avg_launch_ms"], "kernels": kern, "node_bytes": node_b, "bvh_width": s2["bvh_width"],
                            "nodes_per_ray": (s2["nodes_visited"] + s2["nodes_visited_shadow"]) / n_rays, "tris_per_ray": (s2["tris_tested"] + s2["tris_tested_shadow"]) / n_rays,
                            "stage_ms": {"trace": ms_tr, "shadow": ms_sh, "shade": ms_shd, "total": s2["ms_total"]},
                            "note": "durations: CUDA events around every launch in a TIME_STAGES context on the context's own stream; node/triangle "
                                    "counts: a COUNT_TRAVERSAL context on the same steps; k_shade = the per-material shading kernels of one iteration incl. the 2 sort kernels"}
        # ---- CPU baseline (bounded sample): the reference arm's result when it ran on this box just before, else a 1-step run ----
        if not args.no_cpu:
            info = None
            try:
                with open(REF_CACHE) as fh:
                    c = json.load(fh)
                if (c["tris"], c["width"], c["height"], c["ref_spp"], c.get("ref_crop")) == (args.tris, args.width, args.height, args.ref_spp, bool(args.ref_crop)) and time.time() - os.path.getmtime(REF_CACHE) < 6 * 3600:
                    info = c["info"]
                    info["from"] = "the --impl reference run of this bench on this box (%s)" % REF_CACHE
            except Exception:
                pass
            if info is None:
                info = run_cpu_reference(args, 0, 1)
            if info:
                rps, rps_src = rays_per_sample("crop" if args.ref_crop else "frame_1spp")
                line["cpu_baseline"] = cpu_baseline_dict(args, info, rps, rps_src)
                if "from" in info:
                    line["cpu_baseline"]["from"] = info["from"]
        print(json.dumps(line))
    if ctx is not None:
        ctx.close()
    from libyafaray_b200 import unpin_host_arrays
    unpin_host_arrays(pinned)
    if world > 1:
        dist.destroy_process_group()


def C_sizeof(t):
    import ctypes
    return ctypes.sizeof(t)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--spp-per-step", type=int, default=FRAME_SPP)
    ap.add_argument("--ref-spp", type=int, default=1, help="spp of the bounded CPU sample")
    ap.add_argument("--width", type=int, default=WIDTH)
    ap.add_argument("--height", type=int, default=HEIGHT)
    ap.add_argument("--tris", type=int, default=N_TRIS)
    ap.add_argument("--ref-crop", action="store_true", help="CPU sample = centred half-size window instead of the full frame")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--pageable", action="store_true", help="e2e leg: leave the scene arrays pageable (A/B against the default, yafgpu_host_register once before the timed steps)")
    ap.add_argument("--option", action="append", default=[], metavar="NAME=INT", help="yafgpu_set_option knob for A/B runs (e.g. shade_streams=1)")
    ap.add_argument("--wave-size", type=int, default=0, help="camera samples per wavefront wave (0 = library default)")
    args = ap.parse_args()
    rank, world, local_rank = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        reference_arm(args, rank, world)
        return
    if world == 1 and args.gpus > 1:
        # convenience: re-launch under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus), "--master-addr", "127.0.0.1",
               "--master-port", "29571", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    gpu_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
