"""Parity cases shared by tests/ and tools/make_golden.py: name -> (scene builder, kind).

kind "exact"     : deterministic in the reference (no Russian roulette draw, one light): compared per pixel at 1e-5.
kind "converged" : Monte-Carlo; the golden file holds a high-spp reference render (tools/make_golden.py) and the test
                   compares a high-spp GPU render with it (north_star: per-pixel RMSE <= 1e-3 relative to the converged image).
"""
from libyafaray_b200 import scenes

W, H = 64, 36


def _cornell(**kw):
    return lambda spp: scenes.cornell_box(W, H, spp, **kw)


EXACT = {
    "simple_dl_gauss": lambda: scenes.simple_point_scene(160, 90, 1),
    "simple_dl_box4": lambda: scenes.simple_point_scene(160, 90, 4, filter_type="box"),
    "simple_pt_box8": lambda: scenes.simple_point_scene(160, 90, 8, integrator="pathtracing", filter_type="box"),
    "cornell_dl_noglass": lambda: scenes.cornell_box(160, 90, 4, integrator="directlighting", glass=False),
    "cornell_dl_glass": lambda: scenes.cornell_box(160, 90, 4, integrator="directlighting"),
    "cornell_dl_mitchell": lambda: scenes.cornell_box(96, 54, 2, integrator="directlighting", filter_type="mitchell", aa_passes=2),
}

# name -> (builder(spp), converged spp of the golden reference)
CONVERGED = {
    "cornell_pt": (_cornell(path_samples=1, bounces=8, raydepth=8, rr_min_bounces=2), 65536),
    "cornell_pt_noglass": (_cornell(glass=False, path_samples=1, bounces=4, raydepth=4, rr_min_bounces=1), 32768),
    # same scene as cornell_pt at 8x the samples (reference: 86 CPU-minutes on 6 threads, /tmp script = tools/converged_dump.py)
    "cornell_pt_512k": (_cornell(path_samples=1, bounces=8, raydepth=8, rr_min_bounces=2), 524288),
}


import os as _os
ENV_HDR = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "golden", "env_64x32.hdr")


def trace_rays(n, lo, hi, seed=1, tmin=5e-5):
    """Fixed ray set: origins uniform in the (enlarged) scene bound, isotropic directions, seeded."""
    import numpy as np
    rng = np.random.default_rng(seed)
    o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = np.zeros((n, 8), np.float32)
    r[:, 0:3], r[:, 3], r[:, 4:7], r[:, 7] = o, tmin, d, -1.0
    return r


def grid_mesh_scene(n_side=64, seed=3, width=64, height=36):
    """Procedural displaced grid (2*n_side^2 triangles) in the spirit of config 3, small enough for the oracle."""
    import numpy as np
    from libyafaray_b200.interface import Interface
    rng = np.random.default_rng(seed)
    y = Interface()
    y.startScene(0)
    for k, col in enumerate([(0.8, 0.3, 0.3), (0.3, 0.8, 0.3), (0.3, 0.3, 0.8)]):
        y.paramsClearAll(); y.paramsSetString("type", "shinydiffusemat"); y.paramsSetColor("color", *col)
        y.createMaterial("m%d" % k)
    xs = np.linspace(-4, 4, n_side + 1, dtype=np.float32)
    zz = (rng.random((n_side + 1, n_side + 1)) * 0.6).astype(np.float32)
    y.startGeometry()
    y.startTriMesh(1, (n_side + 1) ** 2, 2 * n_side * n_side, False, False, 0)
    for j in range(n_side + 1):
        for i in range(n_side + 1):
            y.addVertex(float(xs[i]), float(xs[j]), float(zz[j, i]))
    for j in range(n_side):
        for i in range(n_side):
            a = j * (n_side + 1) + i
            m = "m%d" % ((i + j) % 3)
            y.addTriangle(a, a + 1, a + n_side + 2, m)
            y.addTriangle(a, a + n_side + 2, a + n_side + 1, m)
    y.endTriMesh()
    y.endGeometry()
    y.paramsClearAll(); y.paramsSetString("type", "pointlight"); y.paramsSetPoint("from", 0.5, -1.0, 6.0)
    y.paramsSetColor("color", 1, 1, 1); y.paramsSetFloat("power", 40.0); y.createLight("L")
    y.paramsClearAll(); y.paramsSetString("type", "perspective")
    y.paramsSetPoint("from", 0, -9, 5); y.paramsSetPoint("to", 0, 0, 0); y.paramsSetPoint("up", 0, -9, 6)
    y.paramsSetInt("resx", width); y.paramsSetInt("resy", height); y.paramsSetFloat("focal", 1.2); y.createCamera("cam")
    y.paramsClearAll(); y.paramsSetString("type", "constant"); y.paramsSetColor("color", 0.1, 0.1, 0.2); y.createBackground("bg")
    y.paramsClearAll(); y.paramsSetString("type", "directlighting"); y.paramsSetInt("raydepth", 2); y.createIntegrator("pt")
    scenes._vol_none(y)
    scenes.render_block(y, width, height, 1)
    return y


TRACE = {
    "simple": lambda: scenes.simple_point_scene(),
    "cornell": lambda: scenes.cornell_box(64, 36, 1),
    "grid64": lambda: grid_mesh_scene(64),
}
TRACE_BOUNDS = {"simple": ((-14, -14, -4), (14, 14, 6)), "cornell": ((-1.4, -1.4, -0.4), (1.4, 1.4, 2.4)), "grid64": ((-5, -5, -1), (5, 5, 2))}


def _set_params(y, kw):
    for k, v in kw.items():
        if isinstance(v, bool): y.paramsSetBool(k, v)
        elif isinstance(v, int): y.paramsSetInt(k, v)
        elif isinstance(v, float): y.paramsSetFloat(k, v)
        elif isinstance(v, str): y.paramsSetString(k, v)
        elif len(v) == 3: y.paramsSetPoint(k, *v)
        else: y.paramsSetColor(k, *v)


def material_zoo_scene(width=64, height=36, spp=4, integrator="pathtracing", ibl=True, sun=True, meshlight=True, transp_shad=False, only_mat=None,
                       hdr_env=None, extra_lights=(), camera=None, gradient=None, **integ):
    """Every material and light type of the hot path in one small scene (configs 3/4 in miniature): glossy, coated_glossy,
    rough_glass, blend_mat, mirror, glass and shinydiffuse boxes on a ground quad; sun + mesh light + constant-background IBL."""
    from libyafaray_b200.interface import Interface
    y = Interface(); y.startScene(0)

    def mat(name, typ, **kw):
        y.paramsClearAll(); y.paramsSetString("type", typ)
        for k, v in kw.items():
            if isinstance(v, bool): y.paramsSetBool(k, v)
            elif isinstance(v, int): y.paramsSetInt(k, v)
            elif isinstance(v, float): y.paramsSetFloat(k, v)
            elif isinstance(v, str): y.paramsSetString(k, v)
            else: y.paramsSetColor(k, *v)
        return y.createMaterial(name)
    ground = mat("ground", "shinydiffusemat", color=(0.7, 0.7, 0.65))
    mats = [
        mat("gloss", "glossy", color=(0.9, 0.6, 0.2), diffuse_color=(0.5, 0.3, 0.1), diffuse_reflect=0.4, glossy_reflect=0.8, exponent=200.0, as_diffuse=False),
        mat("coated", "coated_glossy", color=(0.2, 0.5, 0.9), diffuse_color=(0.1, 0.2, 0.5), diffuse_reflect=0.5, glossy_reflect=0.7, exponent=80.0, IOR=1.8, as_diffuse=False),
        mat("rough", "rough_glass", IOR=1.5, alpha=0.3, filter_color=(0.9, 1.0, 0.9), transmit_filter=1.0),
        mat("aniso", "glossy", color=(0.8, 0.8, 0.8), diffuse_color=(0.3, 0.3, 0.3), diffuse_reflect=0.3, anisotropic=True, exp_u=20.0, exp_v=800.0, as_diffuse=True),
        mat("mirr", "shinydiffusemat", color=(0.6, 0.2, 0.2), specular_reflect=0.5, mirror_color=(1.0, 1.0, 1.0), fresnel_effect=True, IOR=1.6, diffuse_reflect=0.9),
        mat("glas", "glass", IOR=1.45, filter_color=(1.0, 0.9, 0.9), transmit_filter=0.8, fake_shadows=transp_shad),
        mat("sd_on", "shinydiffusemat", color=(0.3, 0.7, 0.3), diffuse_brdf="oren_nayar", sigma=0.4, transparency=0.3, translucency=0.2),
    ]
    y.paramsClearAll(); y.paramsSetString("type", "blend_mat"); y.paramsSetString("material1", "gloss"); y.paramsSetString("material2", "mirr")
    y.paramsSetFloat("blend_value", 0.35)
    mats.append(y.createMaterial("blend"))
    lm = mat("lightmat", "light_mat", color=(1.0, 0.95, 0.8), power=6.0)
    y.startGeometry()
    y.startTriMesh(1, 4, 2, False, False, 0)
    scenes._quad(y, (-8, -8, 0), (8, -8, 0), (8, 8, 0), (-8, 8, 0), ground)
    y.endTriMesh()
    for i, m in enumerate(mats):
        if only_mat is not None:
            m = mats[only_mat]
        y.startTriMesh(2 + i, 8, 12, False, False, 0)
        x0, y0 = -3.6 + 1.9 * (i % 4), -0.4 + 2.2 * (i // 4)
        scenes._box(y, (x0, y0, 0.003), (x0 + 1.2, y0 + 1.2, 1.0 + 0.15 * i), m, rot_deg=11.0 * i)
        y.endTriMesh()
    if meshlight:
        y.startTriMesh(30, 4, 2, False, False, 0)
        scenes._quad(y, (-1.0, 0.5, 3.5), (-1.0, 1.5, 3.5), (1.0, 1.5, 3.5), (1.0, 0.5, 3.5), lm)   # normal -z
        y.endTriMesh()
    y.endGeometry()
    if meshlight:
        y.paramsClearAll(); y.paramsSetString("type", "meshlight"); y.paramsSetInt("object", 30); y.paramsSetColor("color", 1.0, 0.95, 0.8)
        y.paramsSetFloat("power", 6.0); y.paramsSetInt("samples", 2); y.createLight("ml")
    if sun:
        y.paramsClearAll(); y.paramsSetString("type", "sunlight"); y.paramsSetPoint("direction", 0.4, -0.5, 0.8); y.paramsSetFloat("angle", 2.0)
        y.paramsSetColor("color", 1.0, 0.9, 0.8); y.paramsSetFloat("power", 1.2); y.paramsSetInt("samples", 2); y.createLight("sun")
    for i, lp in enumerate(extra_lights):  # dicts of light parameters; 3-tuples are points, 4-tuples colours
        y.paramsClearAll(); _set_params(y, lp); y.createLight("extra%d" % i)
    y.paramsClearAll(); y.paramsSetString("type", "perspective")
    y.paramsSetPoint("from", 0.5, -7.5, 4.0); y.paramsSetPoint("to", 0.0, 0.8, 0.6); y.paramsSetPoint("up", 0.5, -7.5, 5.0)
    y.paramsSetInt("resx", width); y.paramsSetInt("resy", height); y.paramsSetFloat("focal", 1.3)
    if camera:
        _set_params(y, camera)
    y.createCamera("cam")
    if gradient is not None:
        y.paramsClearAll(); y.paramsSetString("type", "gradientback"); _set_params(y, gradient)
    elif hdr_env:
        # equirectangular HDR environment ("textureback" + image texture): hdr_env = {"mapping": ..., "rotation": ..., "interpolate": ...}
        y.paramsClearAll(); y.paramsSetString("type", "image"); y.paramsSetString("filename", ENV_HDR)
        y.paramsSetString("interpolate", hdr_env.get("interpolate", "bilinear")); y.createTexture("envtex")
        y.paramsClearAll(); y.paramsSetString("type", "textureback"); y.paramsSetString("texture", "envtex")
        y.paramsSetString("mapping", hdr_env.get("mapping", "sphere")); y.paramsSetFloat("rotation", hdr_env.get("rotation", 0.0))
        y.paramsSetFloat("power", hdr_env.get("power", 0.7))
        if hdr_env.get("clamp"):
            y.paramsSetFloat("ibl_clamp_sampling", hdr_env["clamp"])
    else:
        y.paramsClearAll(); y.paramsSetString("type", "constant"); y.paramsSetColor("color", 0.35, 0.45, 0.7); y.paramsSetFloat("power", 0.8)
    y.paramsSetBool("ibl", ibl); y.paramsSetInt("ibl_samples", 3); y.createBackground("bg")
    y.paramsClearAll(); y.paramsSetString("type", integrator)
    y.paramsSetInt("raydepth", integ.get("raydepth", 4)); y.paramsSetInt("shadowDepth", 3); y.paramsSetBool("transpShad", transp_shad)
    if integrator == "pathtracing":
        y.paramsSetInt("path_samples", integ.get("path_samples", 2)); y.paramsSetInt("bounces", integ.get("bounces", 4))
        y.paramsSetInt("russian_roulette_min_bounces", integ.get("rr_min_bounces", 1))
    if integ.get("do_AO"):
        y.paramsSetBool("do_AO", True); y.paramsSetInt("AO_samples", integ.get("AO_samples", 8)); y.paramsSetFloat("AO_distance", integ.get("AO_distance", 1.0))
        y.paramsSetColor("AO_color", 0.9, 0.95, 1.0)
    y.createIntegrator("pt")
    scenes._vol_none(y)
    scenes.render_block(y, width, height, spp)
    return y


# path tracing without Russian roulette and with ONE light is deterministic in the reference as well
EXACT["cornell_pt_norr"] = lambda: scenes.cornell_box(160, 90, 16, path_samples=2, bounces=6, raydepth=6, rr_min_bounces=99)
EXACT["zoo_pt_mesh_norr"] = lambda: material_zoo_scene(96, 54, 4, sun=False, ibl=False, rr_min_bounces=99)
EXACT["zoo_dl_point_free"] = lambda: material_zoo_scene(96, 54, 2, integrator="directlighting", raydepth=3)
EXACT["zoo_dl_hdr_ibl"] = lambda: material_zoo_scene(96, 54, 2, integrator="directlighting", raydepth=3, sun=False, meshlight=False,
                                                     hdr_env={"rotation": 40.0})
EXACT["zoo_dl_hdr_angular"] = lambda: material_zoo_scene(96, 54, 2, integrator="directlighting", raydepth=2, sun=False, hdr_env={"mapping": "angular", "rotation": 25.0, "interpolate": "none", "clamp": 4.0})
for _bk, _bias, _rot in (("disk1", "uniform", 0.0), ("disk2", "edge", 0.0), ("ring", "uniform", 0.0), ("triangle", "center", 15.0), ("hexagon", "uniform", 40.0)):
    EXACT["cornell_dl_dof_" + _bk] = (lambda bk=_bk, bias=_bias, rot=_rot: scenes.cornell_box(96, 54, 4, integrator="directlighting", lens={
        "aperture": 0.12, "dof_distance": 3.2, "bokeh_type": bk, "bokeh_bias": bias, "bokeh_rotation": rot}))
EXACT["zoo_dl_ao"] = lambda: material_zoo_scene(96, 54, 2, integrator="directlighting", raydepth=3, sun=False, do_AO=True, AO_samples=4, AO_distance=1.5)
EXACT["zoo_dl_transpshad"] = lambda: material_zoo_scene(96, 54, 2, integrator="directlighting", raydepth=3, transp_shad=True)

# §8f rank 2: the remaining light types, camera types and the gradient background
_ZOO_DL = dict(integrator="directlighting", raydepth=3, sun=False, meshlight=False, ibl=False)
EXACT["zoo_dl_directional"] = lambda: material_zoo_scene(96, 54, 2, extra_lights=[
    {"type": "directional", "direction": (0.3, -0.4, 0.85), "color": (1.0, 0.95, 0.9, 1.0), "power": 1.1},
    {"type": "directional", "direction": (-0.5, 0.2, 0.8), "infinite": False, "from": (-1.5, 0.5, 0.0), "radius": 2.2, "color": (0.4, 0.6, 1.0, 1.0), "power": 2.0}], **_ZOO_DL)
EXACT["zoo_dl_spot"] = lambda: material_zoo_scene(96, 54, 2, extra_lights=[
    {"type": "spotlight", "from": (-2.0, -3.0, 5.0), "to": (-0.5, 0.8, 0.0), "cone_angle": 28.0, "blend": 0.4, "color": (1.0, 0.9, 0.7, 1.0), "power": 40.0},
    {"type": "spotlight", "from": (0.3, 0.2, 0.6), "to": (2.0, 1.0, 0.0), "cone_angle": 50.0, "blend": 0.1, "color": (0.5, 0.7, 1.0, 1.0), "power": 1.5}], **_ZOO_DL)
EXACT["zoo_dl_spot_soft"] = lambda: material_zoo_scene(96, 54, 2, extra_lights=[
    {"type": "spotlight", "from": (2.5, -2.0, 4.5), "to": (0.5, 1.0, 0.0), "cone_angle": 35.0, "blend": 0.3, "soft_shadows": True, "shadowFuzzyness": 0.8,
     "samples": 3, "color": (1.0, 1.0, 0.9, 1.0), "power": 35.0}], **_ZOO_DL)
EXACT["zoo_dl_sphere"] = lambda: material_zoo_scene(96, 54, 2, extra_lights=[
    {"type": "spherelight", "from": (0.5, 0.0, 3.0), "radius": 0.45, "samples": 3, "color": (1.0, 0.9, 0.8, 1.0), "power": 12.0}], **_ZOO_DL)
EXACT["zoo_pt_sphere_norr"] = lambda: material_zoo_scene(96, 54, 4, sun=False, ibl=False, meshlight=False, rr_min_bounces=99, extra_lights=[
    {"type": "spherelight", "from": (0.5, 0.0, 3.0), "radius": 0.45, "samples": 2, "color": (1.0, 0.9, 0.8, 1.0), "power": 12.0}])
EXACT["zoo_dl_gradient_ibl"] = lambda: material_zoo_scene(96, 54, 2, integrator="directlighting", raydepth=3, sun=False, meshlight=False, ibl=True, gradient={
    "horizon_color": (0.9, 0.8, 0.7, 1.0), "zenith_color": (0.2, 0.4, 0.9, 1.0), "horizon_ground_color": (0.3, 0.25, 0.2, 1.0),
    "zenith_ground_color": (0.05, 0.05, 0.04, 1.0), "power": 1.3})
_CAM_LIGHT = dict(integrator="directlighting", raydepth=3, ibl=False)
EXACT["zoo_dl_cam_ortho"] = lambda: material_zoo_scene(96, 54, 2, camera={"type": "orthographic", "scale": 9.5}, **_CAM_LIGHT)
EXACT["zoo_dl_cam_architect"] = lambda: material_zoo_scene(96, 54, 2, camera={"type": "architect", "focal": 1.1, "aperture": 0.05, "dof_distance": 8.0,
                                                                             "bokeh_type": "pentagon", "bokeh_rotation": 10.0}, **_CAM_LIGHT)
EXACT["zoo_dl_cam_angular"] = lambda: material_zoo_scene(96, 54, 2, camera={"type": "angular", "angle": 55.0, "max_angle": 50.0, "circular": True}, **_CAM_LIGHT)
EXACT["zoo_dl_cam_angular_stereo"] = lambda: material_zoo_scene(96, 54, 2, camera={"type": "angular", "angle": 70.0, "circular": False, "mirrored": True,
                                                                                  "projection": "stereographic"}, **_CAM_LIGHT)
EXACT["zoo_dl_cam_equirect"] = lambda: material_zoo_scene(96, 48, 2, camera={"type": "equirectangular"}, **_CAM_LIGHT)

# rendered only through the reference front end with the GPU integrator linked in (oracle/_ref/yaf_xml_gpu): adaptive AA is
# driven by the reference's own ImageFilm::nextPass there
FRONTEND_ONLY = {
    "cornell_dl_adaptive": lambda: scenes.cornell_box(160, 90, 2, integrator="directlighting", aa_passes=3, AA_threshold=0.02, AA_inc_samples=2),
}
CONVERGED["zoo_pt"] = (lambda spp: material_zoo_scene(W, H, spp), 8192)
