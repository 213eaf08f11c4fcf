"""TEST INFRASTRUCTURE: access to the oracle (the reference's own CPU code built by oracle/Makefile into oracle/_ref/)
and to the golden fixtures generated from it (tests/golden/, written by tools/make_golden.py).

Golden files are preferred; when one is missing and the oracle binary is present (this container, or the GPU box, where
oracle/_ref travels with the snapshot) the oracle is executed instead.  Nothing here is imported by the product package.
"""
import json
import os
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
HIT_DTYPE = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("tri", "<i4")])


def ref_bin(flavour="ieee"):
    p = os.path.join(ROOT, "oracle", "_ref", "yaf_ref_" + flavour)
    return p if os.path.exists(p) else None


def _run(args, timeout=3600, env=None):
    r = subprocess.run(args, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=timeout, env=env)
    if r.returncode != 0:
        raise RuntimeError("oracle failed: %s\n%s" % (" ".join(args), r.stderr.decode(errors="replace")[-2000:]))
    out = r.stdout.decode(errors="replace").strip().splitlines()
    for line in reversed(out):
        if line.startswith("{"):
            return json.loads(line)
    return {}


def load_film(path):
    with open(path, "rb") as fh:
        assert fh.read(4) == b"YGF1"
        w, h = np.frombuffer(fh.read(8), "<i4")
        return np.frombuffer(fh.read(), np.float32).reshape(h, w, 5).copy()


def run_reference_render(iface, threads=1, flavour="ieee", timeout=3600, film_path=None):
    """Render `iface` with the reference CPU code; returns (film[h,w,5], info dict).  `film_path` = the image output path that the
    <render> parameter film_save_load refers to (YAF_REF_FILM_PATH of oracle/ref_driver/yaf_ref.cc)."""
    exe = ref_bin(flavour)
    if exe is None:
        return None, None
    with tempfile.TemporaryDirectory() as td:
        xml, out = os.path.join(td, "scene.xml"), os.path.join(td, "out.film")
        iface.to_xml(xml)
        info = _run([exe, "render", xml, out, str(threads)], timeout, env=None if film_path is None else dict(os.environ, YAF_REF_FILM_PATH=film_path))
        return load_film(out), info


def gpu_frontend_bin():
    """The reference front end with the GPU integrator linked in (integration/gpu_integrator.cc, oracle/Makefile `gpu`)."""
    p = os.path.join(ROOT, "oracle", "_ref", "yaf_xml_gpu")
    return p if os.path.exists(p) else None


def run_gpu_frontend(iface, env=None, timeout=600):
    """XML -> reference loader -> GpuIntegrator -> C ABI -> GPU -> reference ImageFilm; returns (film, info, completed process)."""
    exe = gpu_frontend_bin()
    with tempfile.TemporaryDirectory() as td:
        xml, out = os.path.join(td, "scene.xml"), os.path.join(td, "out.film")
        iface.to_xml(xml)
        e = dict(os.environ)
        e.update(env or {})
        r = subprocess.run([exe, "render", xml, out, "1"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=timeout, env=e)
        film, info = None, {}
        if r.returncode == 0:
            film = load_film(out)
            for line in reversed(r.stdout.decode(errors="replace").strip().splitlines()):
                if line.startswith("{"):
                    info = json.loads(line); break
        return film, info, r


def reference_film(iface, key, allow_run=True, write=False):
    """Accumulator film of the reference for `iface`: golden fixture tests/golden/<key>.npz, else the live oracle."""
    path = os.path.join(GOLDEN, key + ".npz")
    if os.path.exists(path) and not write:
        return np.load(path)["film"]
    if not allow_run:
        return None
    film, _ = run_reference_render(iface)
    if film is not None and write:
        np.savez_compressed(path, film=film)
    return film


def reference_hits(iface, rays, mode=0, shadow_depth=2, key=None, write=False):
    """Scene::intersect / isShadowed of the reference on `rays` [n,8]; golden fixture or live oracle."""
    path = os.path.join(GOLDEN, key + ".npz") if key else None
    if path and os.path.exists(path) and not write:
        z = np.load(path)
        if z["rays"].shape == rays.shape and np.array_equal(z["rays"], rays):
            return z["hits"].view(HIT_DTYPE).reshape(-1)
    exe = ref_bin("ieee")
    if exe is None:
        return None
    with tempfile.TemporaryDirectory() as td:
        xml, rb, hb = os.path.join(td, "scene.xml"), os.path.join(td, "rays.bin"), os.path.join(td, "hits.bin")
        iface.to_xml(xml)
        np.ascontiguousarray(rays, np.float32).tofile(rb)
        _run([exe, "hits", xml, rb, hb, str(mode), str(shadow_depth)])
        hits = np.fromfile(hb, HIT_DTYPE)
    if path and write:
        np.savez_compressed(path, rays=rays, hits=hits.view(np.uint8).reshape(len(hits), 16))
    return hits


def read_kat(path):
    """Records written by `yaf_ref kat` (oracle/ref_driver/kat.h)."""
    out = {}
    dts = {0: "<f4", 1: "<f8", 2: "<i4", 3: "<u4"}
    with open(path, "rb") as fh:
        data = fh.read()
    o = 0
    while o < len(data):
        nl = int(np.frombuffer(data, "<u4", 1, o)[0]); o += 4
        name = data[o:o + nl].decode(); o += nl
        dt = int(np.frombuffer(data, "<u4", 1, o)[0]); o += 4
        cnt = int(np.frombuffer(data, "<u8", 1, o)[0]); o += 8
        arr = np.frombuffer(data, dts[dt], cnt, o).copy(); o += arr.nbytes
        out[name] = arr
    return out


def rel_rmse(img, ref):
    """Per-pixel RMSE over RGB relative to the mean luminance of the reference image."""
    d = (img[..., :3].astype(np.float64) - ref[..., :3].astype(np.float64))
    return float(np.sqrt(np.mean(d * d)) / max(1e-12, float(np.mean(ref[..., :3]))))
