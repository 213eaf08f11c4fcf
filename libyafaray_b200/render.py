"""Host mirror of the reference's render control for the hot path.

`Context` wraps one yafgpu context (one GPU).  `render_scene` follows TiledIntegrator::render
(/root/reference/src/integrator/integrator_tiled.cc:116-259): pass 0 with AA_minsamples, then AA_passes-1 passes of
ceil(AA_inc_samples * multiplier) samples with growing sampling offsets.  Adaptive resampling (ImageFilm::nextPass
flagging) is evaluated on the host film between passes when AA_threshold > 0.
"""
import ctypes as C
import math
import os
import struct

import numpy as np

from . import capi


def _check(ctx, rc):
    if rc != capi.OK:
        msg = capi.lib.yafgpu_last_error(ctx)
        raise capi.YafGpuError(rc, msg.decode() if msg else "?")


class Context:
    """One GPU context of libyafgpu.so (include/yafgpu.h)."""

    def __init__(self, device=0, flags=0):
        h = C.c_void_p()
        rc = capi.lib.yafgpu_create(C.byref(h), device, flags)
        if rc != capi.OK:
            msg = capi.lib.yafgpu_last_error(None)
            raise capi.YafGpuError(rc, msg.decode() if msg else "?")
        self.h = h
        self.width = self.height = 0
        self._keep = None

    def close(self):
        if self.h:
            capi.lib.yafgpu_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, desc, keep=None):
        self._keep = keep
        _check(self.h, capi.lib.yafgpu_upload_scene(self.h, C.byref(desc)))
        self.width, self.height = desc.film.width, desc.film.height

    def scene_bound(self):
        b = (capi.f32 * 6)()
        _check(self.h, capi.lib.yafgpu_scene_bound(self.h, b))
        return np.array(b[:], np.float32)

    def trace(self, rays, mode=capi.TRACE_CLOSEST, shadow_depth=2):
        """rays: float32 [n,8] (from xyz, tmin, dir xyz, tmax). Returns structured hits (t,u,v,tri)."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        hits = np.zeros(len(rays), dtype=[("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("tri", "<i4")])
        _check(self.h, capi.lib.yafgpu_trace(self.h, rays.ctypes.data, len(rays), hits.ctypes.data, mode, shadow_depth))
        return hits

    def trace_static(self, rays, mode=capi.TRACE_CLOSEST):
        """yafgpu_trace_static: the same trace without any device allocation (static scratch, at most capi.TRACE_STATIC_MAX_RAYS rays)."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        hits = np.zeros(len(rays), dtype=[("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("tri", "<i4")])
        _check(self.h, capi.lib.yafgpu_trace_static(self.h, rays.ctypes.data, len(rays), hits.ctypes.data, mode))
        return hits

    def trace_device(self, d_rays_ptr, n, d_hits_ptr, mode=capi.TRACE_CLOSEST, shadow_depth=2):
        _check(self.h, capi.lib.yafgpu_trace_device(self.h, d_rays_ptr, n, d_hits_ptr, mode, shadow_depth))

    def film_clear(self):
        _check(self.h, capi.lib.yafgpu_film_clear(self.h))

    def render_pass(self, n_samples, offset=0, adaptive=False, aa_pass_number=0, mask=None, sample_begin=0, sample_end=0,
                    part_index=0, part_count=1):
        p = capi.PassDesc()
        p.n_samples, p.offset, p.adaptive, p.aa_pass_number = n_samples, offset, int(adaptive), aa_pass_number
        p.sample_begin, p.sample_end, p.part_index, p.part_count = sample_begin, sample_end, part_index, part_count
        if mask is not None:
            mask = np.ascontiguousarray(mask, np.uint8)
            p.resample_mask = mask.ctypes.data_as(C.POINTER(C.c_uint8))
        _check(self.h, capi.lib.yafgpu_render_pass(self.h, C.byref(p)))

    def next_pass(self, threshold, dark_detection_type=0, dark_threshold_factor=0.0, variance_edge_size=10, variance_pixels=0, detect_color_noise=False):
        """ImageFilm::nextPass's resampling decision on the device; returns the number of pixels to resample.  Follow with render_pass(adaptive=2)."""
        prm = capi.AaNoiseParams(threshold, int(dark_detection_type), dark_threshold_factor, int(variance_edge_size), int(variance_pixels), int(detect_color_noise))
        n = C.c_int64(0)
        _check(self.h, capi.lib.yafgpu_next_pass(self.h, C.byref(prm), C.byref(n)))
        return n.value

    def sync(self):
        _check(self.h, capi.lib.yafgpu_sync(self.h))

    def read_film(self):
        out = np.empty((self.height, self.width, 5), np.float32)
        _check(self.h, capi.lib.yafgpu_read_film(self.h, out.ctypes.data))
        return out

    def write_film(self, film):
        film = np.ascontiguousarray(film, np.float32)
        _check(self.h, capi.lib.yafgpu_write_film(self.h, film.ctypes.data))

    def film_device_ptr(self):
        p, n = C.c_void_p(), C.c_size_t()
        _check(self.h, capi.lib.yafgpu_film_device_ptr(self.h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def set_aa_multipliers(self, light, indirect):
        _check(self.h, capi.lib.yafgpu_set_aa_multipliers(self.h, light, indirect))

    def set_option(self, name, value):
        _check(self.h, capi.lib.yafgpu_set_option(self.h, name.encode(), int(value)))

    def stats(self):
        s = capi.Stats()
        _check(self.h, capi.lib.yafgpu_get_stats(self.h, C.byref(s)))
        return {k: (list(getattr(s, k)) if hasattr(getattr(s, k), "__len__") else getattr(s, k)) for k, _ in capi.Stats._fields_}

    def reset_stats(self):
        _check(self.h, capi.lib.yafgpu_reset_stats(self.h))

    def stream(self):
        return capi.lib.yafgpu_stream(self.h)


class Group:
    """N GPUs of one box rendering one scene (yafgpu_group, include/yafgpu.h): the scene is replicated, every pass is split by
    interleaved tiles, `read_film` returns the summed film.  Same methods as `Context`, so `render_scene(..., ctx=Group([0, 1]))` works."""

    def __init__(self, devices=None, flags=0):
        h = C.c_void_p()
        n = 0 if devices is None else len(devices)
        arr = (C.c_int * max(1, n))(*(devices or [0]))
        rc = capi.lib.yafgpu_group_create(C.byref(h), arr if n else None, n, flags)
        if rc != capi.OK:
            msg = capi.lib.yafgpu_group_last_error(None)
            raise capi.YafGpuError(rc, msg.decode() if msg else "?")
        self.h = h
        self.width = self.height = 0
        self._keep = None

    def _check(self, rc):
        if rc != capi.OK:
            msg = capi.lib.yafgpu_group_last_error(self.h)
            raise capi.YafGpuError(rc, msg.decode() if msg else "?")

    def size(self):
        return capi.lib.yafgpu_group_size(self.h)

    def close(self):
        if self.h:
            capi.lib.yafgpu_group_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, desc, keep=None):
        self._keep = keep
        self._check(capi.lib.yafgpu_group_upload_scene(self.h, C.byref(desc)))
        self.width, self.height = desc.film.width, desc.film.height

    def film_clear(self):
        self._check(capi.lib.yafgpu_group_film_clear(self.h))

    def render_pass(self, n_samples, offset=0, adaptive=False, aa_pass_number=0, mask=None, sample_begin=0, sample_end=0, part_index=0, part_count=1):
        p = capi.PassDesc()
        p.n_samples, p.offset, p.adaptive, p.aa_pass_number = n_samples, offset, int(adaptive), aa_pass_number
        p.sample_begin, p.sample_end = sample_begin, sample_end   # the partition fields belong to the group
        if mask is not None:
            mask = np.ascontiguousarray(mask, np.uint8)
            p.resample_mask = mask.ctypes.data_as(C.POINTER(C.c_uint8))
        self._check(capi.lib.yafgpu_group_render_pass(self.h, C.byref(p)))

    def read_film(self):
        out = np.empty((self.height, self.width, 5), np.float32)
        self._check(capi.lib.yafgpu_group_read_film(self.h, out.ctypes.data))
        return out

    def next_pass(self, threshold, dark_detection_type=0, dark_threshold_factor=0.0, variance_edge_size=10, variance_pixels=0, detect_color_noise=False):
        prm = capi.AaNoiseParams(threshold, int(dark_detection_type), dark_threshold_factor, int(variance_edge_size), int(variance_pixels), int(detect_color_noise))
        n = C.c_int64(0)
        self._check(capi.lib.yafgpu_group_next_pass(self.h, C.byref(prm), C.byref(n)))
        return n.value

    def read_film_into(self, ptr):
        self._check(capi.lib.yafgpu_group_read_film(self.h, ptr))

    def sync(self):
        self._check(capi.lib.yafgpu_group_sync(self.h))

    def write_film(self, film):
        film = np.ascontiguousarray(film, np.float32)
        self._check(capi.lib.yafgpu_group_write_film(self.h, film.ctypes.data))

    def set_aa_multipliers(self, light, indirect):
        self._check(capi.lib.yafgpu_group_set_aa_multipliers(self.h, light, indirect))

    def set_option(self, name, value):
        self._check(capi.lib.yafgpu_group_set_option(self.h, name.encode(), int(value)))

    def stats(self):
        s = capi.Stats()
        self._check(capi.lib.yafgpu_group_get_stats(self.h, C.byref(s)))
        return {k: (list(getattr(s, k)) if hasattr(getattr(s, k), "__len__") else getattr(s, k)) for k, _ in capi.Stats._fields_}

    def reset_stats(self):
        for k in range(self.size()):
            _check(None, capi.lib.yafgpu_reset_stats(capi.lib.yafgpu_group_ctx(self.h, k)))


def pin_host_arrays(keep, min_bytes=1 << 20):
    """Page-lock the numpy arrays of a scene description's keep-alive list (yafgpu_host_register), so that every later upload of
    them is a DMA at PCIe speed.  Returns the registered arrays; hand them to unpin_host_arrays before they are freed."""
    done = []

    def walk(o):
        if isinstance(o, np.ndarray):
            if o.nbytes >= min_bytes and o.flags["C_CONTIGUOUS"]:
                _check(None, capi.lib.yafgpu_host_register(o.ctypes.data, o.nbytes))
                done.append(o)
        elif isinstance(o, (list, tuple)):
            for v in o:
                walk(v)
    walk(keep)
    return done


def unpin_host_arrays(arrays):
    for a in arrays:
        _check(None, capi.lib.yafgpu_host_unregister(a.ctypes.data))


def film_to_image(film):
    """Pixel::normalized (util_image_buffers.h:39-43): col / weight when weight != 0."""
    w = film[..., 4:5]
    return np.where(w != 0, film[..., :4] / np.where(w != 0, w, 1), 0).astype(np.float32)


def load_film(path):
    """Read the raw film dump written by oracle/ref_driver/yaf_ref.cc (`YGF1`, w, h, w*h*5 float32)."""
    with open(path, "rb") as fh:
        if fh.read(4) != b"YGF1":
            raise ValueError("not a YGF1 film")
        w, h = struct.unpack("<ii", fh.read(8))
        return np.frombuffer(fh.read(), np.float32).reshape(h, w, 5).copy()


def render_scene(iface, output=None, device=0, ctx=None, image_output_path=None):
    """TiledIntegrator::render pass schedule (integrator_tiled.cc:116-259) driving the GPU passes. Returns the film.

    `image_output_path` plays the role of yafaray-xml's output image path for the <render> parameter `film_save_load`
    (environment.cc:499,524-526): "save" writes "<path> - node NNNN.film" after the render, "load-save" first sums every matching
    film of that folder into the accumulators and continues from the loaded sampling offset, skipping pass 1
    (imagefilm.cc:262-267,1467-1558; integrator_tiled.cc:174-202).  The files are the reference's YAF_FILMv1 format (film_io.py)."""
    from . import film_io
    g, rp = iface._g, iface.render_params
    desc, keep = iface.scene_desc()
    own = ctx is None
    if own:
        ctx = Context(device)
    ctx.upload(desc, keep)
    aa_passes, aa_samples = int(g(rp, "AA_passes", 1)), int(g(rp, "AA_minsamples", 1))
    aa_inc = int(g(rp, "AA_inc_samples", aa_samples))
    smf = np.float32(g(rp, "AA_sample_multiplier_factor", 1.0))
    lmf = np.float32(g(rp, "AA_light_sample_multiplier_factor", 1.0))
    imf = np.float32(g(rp, "AA_indirect_sample_multiplier_factor", 1.0))
    threshold = np.float32(g(rp, "AA_threshold", 0.05))
    noise_prm = dict(dark_detection_type=capi.DARK_DETECTION.get(g(rp, "AA_dark_detection_type", "none"), 0), dark_threshold_factor=g(rp, "AA_dark_threshold_factor", 0.0),
                     variance_edge_size=int(g(rp, "AA_variance_edge_size", 10)), variance_pixels=int(g(rp, "AA_variance_pixels", 0)),
                     detect_color_noise=bool(g(rp, "AA_detect_color_noise", False)))
    # integrator_tiled.cc:142: the floor below which the threshold is tightened for the following pass
    floor_pixels = int(math.floor(np.float32(np.float32(g(rp, "AA_resampled_floor", 0.0)) * np.float32(desc.film.width * desc.film.height)) / np.float32(100.0)))
    fsl = g(rp, "film_save_load", "none") if image_output_path is not None else "none"
    node, base_raw = int(g(rp, "adv_computer_node", 0)), int(g(rp, "adv_base_sampling_offset", 0))
    loaded = None
    if fsl == "load-save":
        loaded, _paths = film_io.load_films_in_folder(image_output_path, (desc.film.height, desc.film.width))
        if loaded is not None:
            base_raw = max(base_raw, loaded.base_sampling_offset)
    if fsl in ("save", "load-save"):
        fp = film_io.film_path(image_output_path, node)
        if os.path.exists(fp):   # imageFilmFileBackup (imagefilm.cc:1659-1690)
            os.replace(fp, fp + "-previous.bak")
    base = base_raw + node * 100000   # ImageFilm::getBaseSamplingOffset (imagefilm.h:124)
    if loaded is not None:
        ctx.write_film(loaded.film)
        acum = sampling_offset = loaded.sampling_offset
    else:
        ctx.film_clear()
        ctx.render_pass(aa_samples, 0 + base, False, 0)
        acum = sampling_offset = aa_samples
    sm, lm, im = np.float32(1), np.float32(1), np.float32(1)
    threshold_changed, resampled = True, 0
    for i in range(1, aa_passes):
        sm, lm, im = np.float32(sm * smf), np.float32(lm * lmf), np.float32(im * imf)
        ctx.set_aa_multipliers(float(lm), float(im))
        n = int(math.ceil(np.float32(aa_inc) * sm))
        # integrator_tiled.cc:220-251: the film decides which pixels the pass resamples (ImageFilm::nextPass, evaluated on the device: the film stays in HBM);
        # a pass after one that resampled nothing is skipped unless the threshold moved; too few resampled pixels tighten the threshold
        if resampled <= 0 and not threshold_changed:
            pass
        else:
            resampled = ctx.next_pass(float(threshold), **noise_prm)
            threshold_changed = False
        if resampled > 0:
            ctx.render_pass(n, acum + base, 2, i)
            sampling_offset = acum + n            # ImageFilm::setSamplingOffset(offset + samples) (integrator_tiled.cc:270)
        acum += n
        if resampled < floor_pixels:
            ratio = min(np.float32(8.0), np.float32(floor_pixels) / np.float32(resampled)) if resampled else np.float32(8.0)
            threshold = np.float32(threshold * np.float32(np.float32(1.0) - np.float32(0.1) * ratio))
            if threshold > 0:
                threshold_changed = True
    ctx.sync()
    film = ctx.read_film()
    if fsl in ("save", "load-save"):
        film_io.write_film_file(film_io.film_path(image_output_path, node), film, sampling_offset, base_raw, node,
                                int(g(rp, "xstart", 0)), int(g(rp, "ystart", 0)))
    if output is not None:
        output(film)
    if own:
        ctx.close()
    return film
