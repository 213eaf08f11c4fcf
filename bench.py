"""Benchmark of the path-tracing hot path (contract: python bench.py --gpus N --steps K --warmup W [--impl reference]).

Workload (BASELINE.json configs[1]): Cornell box, PathIntegrator, 1920x1080, 256 spp pass, area light, ShinyDiffuse +
Glass, path_samples 1, bounces 8, raydepth 8, russian_roulette_min_bounces 2.  A "step" renders
`--spp-per-step` x N samples per pixel of that 256-spp pass (sample sub-ranges keep every sample's QMC numbers those of
the full pass), tile-interleaved over the N GPUs, scene replicated; with N > 1 every step ends with the NCCL sum of the
film accumulators on rank 0 (the film is a linear accumulator).  Per-GPU work is fixed -> weak scaling.

  value          whole-job Mrays/s (all ray kinds, counted on the device) with the scene resident in HBM
  samples_per_s  camera samples/s (the reference has no ray counter; its Mrays/s is samples/s x the device-counted
                 rays per camera sample of this very workload, labelled derived)
  e2e            same metric through the public call with HOST buffers: scene upload (H2D + BVH build), film clear,
                 render, film read-back (D2H), every step
  roofline       the closest-hit trace kernel: algorithmic bytes (48 B ray+hit, 64 B per node visited,
                 48 B per triangle tested, counted by a COUNT_TRAVERSAL context on the same steps) / its mean duration
  cpu_baseline   the reference's own CPU code (oracle/_ref/yaf_ref_fast, reference Release flags, all host threads) on
                 a bounded sample of the same workload (same scene and resolution, fewer spp)
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

WIDTH, HEIGHT, PASS_SPP = 1920, 1080, 256
RAYS_FILE = os.path.join(ROOT, "profiles", "rays_per_sample.json")


def build_scene(width, height, spp):
    from libyafaray_b200 import scenes
    return scenes.cornell_box(width, height, spp, path_samples=1, bounces=8, raydepth=8, rr_min_bounces=2, light_samples=4,
                              integrator="pathtracing", threads=-1)


def config_dict(args, world):
    return {"workload": "cornell_box_pathtracing_%dx%d_%dspp_pass" % (args.width, args.height, PASS_SPP),
            "integrator": "pathtracing", "path_samples": 1, "bounces": 8, "raydepth": 8, "rr_min_bounces": 2,
            "light": "arealight samples=4", "materials": "shinydiffusemat+glass", "triangles": 36,
            "spp_per_step_per_gpu": args.spp_per_step, "camera_samples_per_step": args.width * args.height * args.spp_per_step * world,
            "partition": "tiles32 interleaved x%d, scene replicated" % world,
            "cache": "working set (wavefront queues, >1 GB per wave) exceeds the 126 MB L2; no explicit flush"}


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


def run_cpu_reference(width, height, spp, threads, timeout=3000):
    """Times oracle/_ref/yaf_ref_fast (the reference's own CPU code) on the workload at `spp`; returns info dict."""
    exe = os.path.join(ROOT, "oracle", "_ref", "yaf_ref_fast")
    if not os.path.exists(exe):
        return None
    y = build_scene(width, height, spp)
    with tempfile.TemporaryDirectory() as td:
        xml, out = os.path.join(td, "scene.xml"), os.path.join(td, "out.film")
        y.to_xml(xml)
        r = subprocess.run([exe, "render", xml, out, str(threads)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, timeout=timeout)
        for line in reversed(r.stdout.decode(errors="replace").splitlines()):
            if line.startswith("{"):
                return json.loads(line)
    return None


def rays_per_sample_default():
    try:
        with open(RAYS_FILE) as fh:
            return float(json.load(fh)["rays_per_camera_sample"])
    except Exception:
        return 11.147  # device-counted rays per camera sample of this workload (profiles/r1_bench_n1.json); the CPU reference has no ray counter


def reference_arm(args, rank, world):
    """--impl reference: the reference's own CPU implementation on the host cores (rank 0 only)."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    spp = args.ref_spp
    rps = rays_per_sample_default()
    times, samples = [], args.width * args.height * spp
    info = None
    for i in range(args.warmup + args.steps):
        info = run_cpu_reference(args.width, args.height, spp, -1)
        if info is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/yaf_ref_fast is not built"}))
            return
        if i >= args.warmup:
            times.append(float(info["rendert_s"]))
    t = sum(times) / len(times)
    sps = samples / t
    val = sps * rps / 1e6
    sample = "%dx%d at %d spp per step (same scene, sampler and integrator settings; the full pass is %d spp)" % (args.width, args.height, spp, PASS_SPP)
    print(json.dumps({
        "impl": "reference", "metric": "Mrays/s", "value": val, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "samples_per_s": sps, "rays_per_camera_sample": rps, "rays_derived": True, "config": config_dict(args, world),
        "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": int(info.get("threads", cores)), "kind": "reference", "sample": sample,
                         "samples_per_s": sps, "note": "reference rendert timer (integrator_tiled.cc:168-169); kd-tree build excluded; global film mutex (imagefilm.cc:960)"},
        "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


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
    y = build_scene(args.width, args.height, PASS_SPP)
    desc, keep = y.scene_desc()
    ctx = Context(local_rank)
    if args.wave_size:
        ctx.set_option("wave_size", args.wave_size)
    for kv in args.option:
        ctx.set_option(kv.split("=")[0], int(kv.split("=")[1]))
    ctx.upload(desc, keep)
    S = args.spp_per_step * world
    n_chunks = max(1, PASS_SPP // S)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=local_rank)
    def wrap_film(c):
        fptr, fcount = c.film_device_ptr()

        class _Film:
            __cuda_array_interface__ = {"shape": (fcount,), "typestr": "<f4", "data": (fptr, False), "version": 3, "strides": None}
        return torch.as_tensor(_Film(), device="cuda:%d" % local_rank)
    films = {}
    red = torch.empty_like(wrap_film(ctx)) if world > 1 else None

    def step(i, c=ctx):
        b = (i % n_chunks) * S
        c.render_pass(PASS_SPP, 0, False, 0, sample_begin=b, sample_end=min(PASS_SPP, b + S), part_index=rank, part_count=world)
        if world > 1:
            if id(c) not in films:
                films[id(c)] = wrap_film(c)
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
    if world > 1:
        torch.cuda.current_stream().synchronize()
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
            "samples_per_s": samples / sec, "rays_per_camera_sample": rays / max(1.0, samples), "config": config_dict(args, world),
            "gpu_launches": int(launches), "clocks": sampler.summary(), "waves_per_step": st["waves"] / max(1, args.steps)}

    # ---- end-to-end through the public call with host buffers (every step: upload + BVH build, clear, render, film D2H) ----
    host_film = torch.empty((args.height, args.width, 5), dtype=torch.float32, pin_memory=True)  # pinned: the D2H of the film runs at PCIe speed
    g = desc.geometry
    h2d = (g.n_points * 12 + g.n_normals * 12 + g.n_uv * 8 + g.n_tris * (12 * 3 + 8) + desc.n_materials * C_sizeof(capi.Material) + desc.n_lights * C_sizeof(capi.Light))
    e2e_steps = max(2, min(args.steps, 8))

    def e2e_step(i):
        ctx.upload(desc, keep)
        films.pop(id(ctx), None)  # the film accumulators were re-allocated by the upload
        ctx.film_clear()
        step(i)
        capi.lib.yafgpu_read_film(ctx.h, host_film.data_ptr())
    e2e_step(0)
    ctx.reset_stats()
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
                   "ms_per_step": t_e2e / e2e_steps * 1e3, "timing": "host wall clock around upload+clear+render+read_film, max over ranks"}

    if rank == 0:
        # ---- roofline of the closest-hit kernel: counting + stage-timing context on the same steps ----
        world1 = world
        step0 = lambda i, c: c.render_pass(PASS_SPP, 0, False, 0, sample_begin=(i % n_chunks) * S, sample_end=min(PASS_SPP, (i % n_chunks) * S + S), part_index=rank, part_count=world1)

        def side_run(flags):
            c2 = Context(local_rank, flags)
            if args.wave_size:
                c2.set_option("wave_size", args.wave_size)
            for kv in args.option:
                c2.set_option(kv.split("=")[0], int(kv.split("=")[1]))
            c2.upload(desc, keep)
            c2.film_clear()
            step0(0, c2)
            c2.reset_stats()
            for i in range(min(args.steps, 4)):
                step0(args.warmup + i, c2)
            r = c2.stats()
            c2.close()
            return r
        s2 = side_run(capi.FLAG_TIME_STAGES)        # kernel durations (CUDA events around every launch, no counting code)
        s3 = side_run(capi.FLAG_COUNT_TRAVERSAL)    # nodes visited / triangles tested on the same steps
        s2["nodes_visited"], s2["tris_tested"] = s3["nodes_visited"], s3["tris_tested"]
        line["wave0_tasks"], line["wave0_shadow"] = s2["last_wave_tasks"], s2["last_wave_shadow"]
        peak, how = hbm_peak()
        n_trace, n_shadow = s2["rays_in_timed_trace"], s2["rays_in_timed_shadow"]
        ms_tr, ms_sh, ms_shd = s2["ms_trace"], s2["ms_shadow"], s2["ms_shade"]
        # Algorithmic bytes.  Traversal: 48 B ray+hit, 64 B per node visited, 48 B per triangle
        # tested (counters cover closest-hit AND shadow rays).  Shading: per task 108 B queue record in (ray_o, ray_d, hit, wgt,
        # thr|dc, st, perm), 104 B per task spawned, 48 B per shadow ray queued, 136 B of triangle/vertex/normal data per hit
        # when the scene does not fit the caches (36 triangles do: counted 0 here).
        node_b, tri_b = 64, 48
        alg_trav = 48.0 * (n_trace + n_shadow) + node_b * s2["nodes_visited"] + tri_b * s2["tris_tested"]
        spawned = max(0, n_trace - s2["camera_samples"])
        geom_b = 136.0 if desc.geometry.n_tris > 100000 else 0.0
        alg_shade = (108.0 + geom_b) * n_trace + 104.0 * spawned + 48.0 * n_shadow
        kern = {
            "k_shade": {"ms": ms_shd, "launches": s2["launches_shade"], "alg_bytes": alg_shade, "units": n_trace, "unit": "task"},
            "k_trace+k_shadow": {"ms": ms_tr + ms_sh, "launches": s2["launches_trace"] + s2["launches_shadow"], "alg_bytes": alg_trav,
                                 "units": n_trace + n_shadow, "unit": "ray"},
        }
        dom = max(kern, key=lambda k: kern[k]["ms"])
        for k, v in kern.items():
            v["achieved_gbs"] = v["alg_bytes"] / (v["ms"] / 1e3) / 1e9 if v["ms"] > 0 else 0.0
            v["frac"] = v["achieved_gbs"] / peak
            v["avg_launch_ms"] = v["ms"] / max(1, v["launches"])
            v["alg_bytes_per_unit"] = v["alg_bytes"] / max(1, v["units"])
            v["share_of_step"] = v["ms"] / max(1e-9, s2["ms_total"])
        traffic = None
        try:   # dram bytes per unit of the dominant kernel from the committed ncu --set full capture (profiles/)
            with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "ncu_traffic.json")) as fh:
                t = json.load(fh).get(dom)
            if t:
                traffic = t["dram_bytes_per_unit"] * kern[dom]["units"] / max(1, kern[dom]["launches"])
        except Exception:
            pass
        d = kern[dom]
        line["roofline"] = {"bound": "hbm", "achieved": d["achieved_gbs"], "peak": peak, "unit": "GB/s", "frac": d["frac"], "traffic": traffic,
                            "peak_source": how, "kernel": dom, "algorithmic_bytes_per_launch": d["alg_bytes"] / max(1, d["launches"]),
                            "avg_launch_ms": d["avg_launch_ms"], "kernels": kern,
                            "nodes_per_ray": s2["nodes_visited"] / max(1, n_trace + n_shadow), "tris_per_ray": s2["tris_tested"] / max(1, n_trace + n_shadow),
                            "stage_ms": {"trace": ms_tr, "shadow": ms_sh, "shade": ms_shd, "total": s2["ms_total"]},
                            "note": "durations: CUDA events around every launch in a TIME_STAGES context on the context's own stream; node/triangle "
                                    "counts: a COUNT_TRAVERSAL context on the same steps; k_shade = the per-material shading kernels of one iteration "
                                    "(incl. the 2 sort kernels); Cornell's 36 triangles live in L1, so its traversal figure is cache bandwidth, not HBM"}
        # ---- CPU baseline (bounded sample) ----
        if not args.no_cpu:
            cores = os.cpu_count() or 1
            info = run_cpu_reference(args.width, args.height, args.ref_spp, -1)
            if info:
                sps = args.width * args.height * args.ref_spp / float(info["rendert_s"])
                rps = line["rays_per_camera_sample"]
                line["cpu_baseline"] = {"value": sps * rps / 1e6, "unit": "Mrays/s", "cores": int(info.get("threads", cores)), "kind": "reference",
                                        "sample": "%dx%d at %d spp, same scene/sampler/integrator (full pass: %d spp); Mrays/s derived = CPU samples/s x device-counted rays per sample" % (args.width, args.height, args.ref_spp, PASS_SPP),
                                        "samples_per_s": sps, "rendert_s": info["rendert_s"]}
        try:
            os.makedirs(os.path.dirname(RAYS_FILE), exist_ok=True)
            if args.width == WIDTH and args.height == HEIGHT and world == 1:
                with open(RAYS_FILE, "w") as fh:
                    json.dump({"rays_per_camera_sample": line["rays_per_camera_sample"], "workload": line["config"]["workload"]}, fh)
        except Exception:
            pass
        print(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def C_sizeof(t):
    import ctypes
    return ctypes.sizeof(t)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--spp-per-step", type=int, default=4)
    ap.add_argument("--ref-spp", type=int, default=1, help="spp of the bounded CPU sample")
    ap.add_argument("--width", type=int, default=WIDTH)
    ap.add_argument("--height", type=int, default=HEIGHT)
    ap.add_argument("--no-cpu", action="store_true")
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
