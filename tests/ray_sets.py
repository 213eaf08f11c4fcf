"""TEST INFRASTRUCTURE: the fixed ray sets of SURVEY.md §8(d) for hit-id parity at 10^6 - 10^7 triangles.

  * n_cam camera rays     : from the scene camera's position through image-plane points drawn from the generator;
  * n_surf secondary rays : origins on uniformly drawn triangles at uniformly drawn barycentric points, lifted off the surface
                            along the face normal, cosine-hemisphere directions about that normal (the shape of the rays a
                            path tracer's diffuse bounce produces).
All random numbers come from the reference's own MWC generator `Random`, seed 1
(/root/reference/include/utility/util_mcqmc.h:173-190), restated in `mwc_stream` (pinned against the reference's known-answer
dump in tests/test_cpu.py).
"""
import hashlib
import os

import numpy as np

MWC_A = 1791398085


def mwc_reference_loop(seed, n):
    """Random::operator() restated literally (16-bit partial products and the carry test), for the known-answer test."""
    x, c = 30903, seed & 0xffffffff
    ah, al = MWC_A >> 16, MWC_A & 65535
    out = np.empty(n, np.float64)
    for i in range(n):
        xh, xl = x >> 16, x & 65535
        nx = (x * MWC_A + c) & 0xffffffff
        c = (xh * ah + (((xh * al) & 0xffffffff) >> 16) + (((xl * ah) & 0xffffffff) >> 16)) & 0xffffffff
        if ((xl * al) & 0xffffffff) >= ((~c + 1) & 0xffffffff):
            c = (c + 1) & 0xffffffff
        x = nx
        out[i] = x * 2.0 ** -32
    return out


def mwc_stream(seed, n):
    """First n outputs of Random(seed) as float64 in [0,1).  The reference's carry arithmetic is NOT an exact multiply-with-carry
    (the 16-bit partial products drop low bits), so there is no closed form to jump ahead with: the sequence is evaluated
    literally, on Python ints, about 1.5 us per draw."""
    x, c = 30903, seed & 0xffffffff
    ah, al, a, m = MWC_A >> 16, MWC_A & 65535, MWC_A, 0xffffffff
    out = [0] * n
    for i in range(n):
        xh, xl = x >> 16, x & 65535
        nc = xh * ah + (((xh * al) & m) >> 16) + (((xl * ah) & m) >> 16)
        x = (x * a + c) & m
        c = nc & m
        if ((xl * al) & m) >= ((-c) & m):
            c = (c + 1) & m
        out[i] = x
    return np.array(out, np.float64) * 2.0 ** -32


def camera_rays(iface, u):
    """Rays from the perspective camera of `iface` through the image-plane points u[n,2] in [0,1)^2 (PerspectiveCamera
    constructor geometry, /root/reference/src/camera/camera_perspective.cc:34-75, in float64; these are probe rays, they do
    not have to equal the integrator's own camera rays bit for bit)."""
    cam = next(iter(iface.cameras.values()))
    g = lambda k, d: cam[k][1] if k in cam else d
    pos, to, up = (np.array(g(k, d), np.float64) for k, d in (("from", (0, 1, 0)), ("to", (0, 0, 0)), ("up", (0, 1, 1))))
    resx, resy, focal = g("resx", 320), g("resy", 200), g("focal", 1.0)
    aspect = g("aspect_ratio", 1.0) * resy / resx
    z = to - pos; z /= np.linalg.norm(z)
    x = np.cross(up - pos, z); x /= np.linalg.norm(x)
    y = np.cross(z, x)
    px, py = (u[:, 0] - 0.5), (u[:, 1] - 0.5) * aspect
    d = px[:, None] * x[None] + py[:, None] * (-y)[None] + focal * z[None]
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = np.zeros((len(u), 8), np.float32)
    r[:, 0:3], r[:, 3], r[:, 4:7], r[:, 7] = pos, 0.0, d, -1.0
    return r


def surface_rays(points, faces, u, tmin=5e-5):
    """u[n,5]: triangle pick, two barycentric draws, two hemisphere draws."""
    n = len(u)
    t = np.minimum((u[:, 0] * len(faces)).astype(np.int64), len(faces) - 1)
    f = faces[t]
    a, b, c = (points[f[:, k]].astype(np.float64) for k in range(3))
    su = np.sqrt(u[:, 1])
    b0, b1 = 1.0 - su, u[:, 2] * su
    p = a * b0[:, None] + b * b1[:, None] + c * (1.0 - b0 - b1)[:, None]
    ng = np.cross(b - a, c - a)
    ln = np.linalg.norm(ng, axis=1, keepdims=True)
    ng = np.where(ln > 0, ng / np.maximum(ln, 1e-300), np.array([0.0, 0.0, 1.0]))
    # cosine hemisphere about ng (util_sample.h:45-60 shape: z = sqrt(1 - s1), phi = 2 pi s2)
    zc = np.sqrt(1.0 - u[:, 3]); rr = np.sqrt(u[:, 3]); phi = 2.0 * np.pi * u[:, 4]
    h = np.where(np.abs(ng[:, 0:1]) > 0.5, np.array([[0.0, 1.0, 0.0]]), np.array([[1.0, 0.0, 0.0]]))
    tu = np.cross(h, ng); tu /= np.linalg.norm(tu, axis=1, keepdims=True)
    tv = np.cross(ng, tu)
    d = tu * (rr * np.cos(phi))[:, None] + tv * (rr * np.sin(phi))[:, None] + ng * zc[:, None]
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = np.zeros((n, 8), np.float32)
    r[:, 0:3] = p + ng * 1e-4
    r[:, 3], r[:, 4:7], r[:, 7] = tmin, d, -1.0
    return r


def scene_triangles(iface):
    """(points, faces) of all meshes in global triangle order (object id order, Scene::update scene.cc:807-817)."""
    P, F, off = [], [], 0
    for oid in sorted(iface.meshes):
        m = iface.meshes[oid]
        if m.np_faces is not None:
            p, f = m.np_points, m.np_faces
        else:
            pts = m.points[0::2] if m.has_orco else m.points
            p, f = np.array(pts, np.float32).reshape(-1, 3), np.array(m.faces, np.int32).reshape(-1, 3)
        P.append(p); F.append(f + off); off += len(p)
    return np.concatenate(P), np.concatenate(F)


def fixed_ray_set(iface, n_cam=1 << 20, n_surf=1 << 20, seed=1):
    u = mwc_stream(seed, 2 * n_cam + 5 * n_surf)
    cam = camera_rays(iface, u[:2 * n_cam].reshape(-1, 2))
    P, F = scene_triangles(iface)
    surf = surface_rays(P, F, u[2 * n_cam:].reshape(-1, 5))
    return np.concatenate([cam, surf])


# ---- the scale scenes (geometry of BASELINE configs 3 and 4) and the cache of reference answers ----------------------------
SCALE = {
    # name: (kwargs of scenes.procedural_mesh, rays per set)
    "c4": (dict(n_tris=1_000_000, mesh_lights=3, blend=True), 1 << 20),
    "c3": (dict(n_tris=10_000_000), 1 << 20),
    "c5": (dict(n_tris=20_000_000, seed=23456), 1 << 20),   # BASELINE configs[4]: the config-3 generator at 2 * 10^7 triangles, seed 23456
}


def scale_scene(name, width=160, height=90, spp=1, integrator="directlighting"):
    from libyafaray_b200 import scenes
    kw, _ = SCALE[name]
    return scenes.procedural_mesh(width=width, height=height, spp=spp, integrator=integrator, **kw)


def cache_dir():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    d = os.path.join(root, "oracle", "_ref", "cache")
    os.makedirs(d, exist_ok=True)
    return d


def rays_digest(rays):
    return hashlib.sha256(np.ascontiguousarray(rays).tobytes()).hexdigest()[:16]


def mwc_cached(seed, n):
    """`mwc_stream`, memoised as uint32 under oracle/_ref/cache (10 s of Python per 7 M draws otherwise)."""
    p = os.path.join(cache_dir(), "mwc_seed%d_%d.npy" % (seed, n))
    if os.path.exists(p):
        return np.load(p).astype(np.float64) * 2.0 ** -32
    u = mwc_stream(seed, n)
    try:
        np.save(p, (u * 2.0 ** 32).astype(np.uint32))
    except OSError:
        pass
    return u


def scale_rays(name, iface=None):
    """The fixed ray set of a scale scene: n camera rays followed by n secondary rays."""
    iface = iface or scale_scene(name)
    n = SCALE[name][1]
    u = mwc_cached(1, 7 * n)
    cam = camera_rays(iface, u[:2 * n].reshape(-1, 2))
    P, F = scene_triangles(iface)
    return np.concatenate([cam, surface_rays(P, F, u[2 * n:].reshape(-1, 5))])


def scale_reference_hits(name, iface, rays, allow_run=True):
    """Scene::intersect of the reference on the scale ray set: the cached answer under oracle/_ref/cache (written by
    tools/make_scale_hits.py in the build container, travels with the snapshot) when it belongs to exactly these rays, else the
    live oracle (XML export + kd-tree build: ~10 s at 10^6 triangles, ~2 min at 10^7)."""
    from tests import oracle_util as ou
    p = os.path.join(cache_dir(), "hits_%s_%s.npy" % (name, rays_digest(rays)))
    if os.path.exists(p):
        return np.load(p).view(ou.HIT_DTYPE).reshape(-1), "cache"
    if not allow_run or ou.ref_bin() is None:
        return None, "unavailable"
    h = ou.reference_hits(iface, rays, 0)
    try:
        np.save(p, h.view(np.uint8).reshape(len(h), 16))
    except OSError:
        pass
    return h, "live"


def triangle_intersect_f32(points, faces, rays, tri):
    """Triangle::intersect (/root/reference/include/common/triangle.h:223-259) restated on float32 numpy arrays with the reference's
    operation order (no fused multiply-add): returns (ok, t, u, v) for ray i against triangle tri[i].  Used to certify that a
    triangle the GPU reports is a hit under the reference's own arithmetic."""
    f = np.float32
    o, d = rays[:, 0:3].astype(f), rays[:, 4:7].astype(f)
    a, b, c = (points[faces[tri][:, k]].astype(f) for k in range(3))
    e1, e2 = b - a, c - a

    def cross(x, y):
        return np.stack([x[:, 1] * y[:, 2] - x[:, 2] * y[:, 1], x[:, 2] * y[:, 0] - x[:, 0] * y[:, 2], x[:, 0] * y[:, 1] - x[:, 1] * y[:, 0]], axis=1)

    def dot(x, y):
        return (x[:, 0] * y[:, 0] + x[:, 1] * y[:, 1]) + x[:, 2] * y[:, 2]

    def length(x):
        return np.sqrt(dot(x, x))
    eps = ((np.float64(f(0.1)) * 0.00005) * np.maximum(length(e1), length(e2)).astype(np.float64)).astype(f)  # triangle.h:206
    pvec = cross(d, e2)
    det = dot(e1, pvec)
    with np.errstate(all="ignore"):
        inv = f(1.0) / det
        tvec = o - a
        u = dot(tvec, pvec) * inv
        qvec = cross(tvec, e1)
        v = dot(d, qvec) * inv
        t = dot(e2, qvec) * inv
    ok = ~((det > -eps) & (det < eps)) & ~((u < 0) | (u > 1)) & ~((v < 0) | ((u + v) > 1)) & ~(t < eps)
    return ok, t, u, v
