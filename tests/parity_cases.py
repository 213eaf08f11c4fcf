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
    # round 2: the filter and clamp corners of ImageFilm::addSample (imagefilm.cc:60-122,925-1015) and the bg_transp alpha channel
    "cornell_dl_lanczos": lambda: scenes.cornell_box(96, 54, 2, integrator="directlighting", filter_type="lanczos"),
    "cornell_dl_lanczos_wide": lambda: scenes.cornell_box(96, 54, 2, integrator="directlighting", filter_type="lanczos", pixelwidth=3.0),
    "cornell_dl_clamp": lambda: scenes.cornell_box(96, 54, 4, integrator="directlighting", AA_clamp_samples=0.7, AA_clamp_indirect=0.3),
    "cornell_pt_clamp_norr": lambda: scenes.cornell_box(96, 54, 4, path_samples=2, bounces=3, raydepth=3, rr_min_bounces=8, AA_clamp_samples=0.5, AA_clamp_indirect=0.2),
    # glass / rough-glass dispersion: the dispersive branch of recursiveRaytrace (integrator_montecarlo.cc:796-858), Cauchy IOR per wavelength
    # (material_glass.cc:64-216,242-309, material_rough_glass.cc:61-172), wl2Rgb__ through the CIE table
    "cornell_dl_dispersion": lambda: scenes.cornell_box(96, 54, 2, integrator="directlighting", dispersion=20.0, raydepth=4),
    "cornell_pt_dispersion_norr": lambda: scenes.cornell_box(96, 54, 2, path_samples=2, bounces=3, raydepth=3, rr_min_bounces=8, dispersion=20.0),
    # rough glass is glossy AND dispersive: its glossy-branch samples read RenderState::wavelength_, which the previous sibling's subtree
    # overwrites in the reference (a leftover of its depth-first order, like include_lights_); exact only while the children cannot recurse
    "cornell_dl_rough_dispersion": lambda: scenes.cornell_box(96, 54, 2, integrator="directlighting", dispersion=30.0, rough_glass=True, raydepth=1),
    "simple_dl_bgtransp": lambda: scenes.simple_point_scene(160, 90, 2, filter_type="box", bg_transp=True),
    # "bg_transp_refract": Material::getAlpha folded along the refraction chain (integrator_path_tracer.cc:319-328, integrator_montecarlo.cc:954,1023);
    # glass, tinted glass, rough glass, transparent shinydiffuse and a blend in front of the open background, with and without a ground plane behind
    "simple_dl_bgtransp_refract": lambda: scenes.simple_point_scene(160, 90, 2, filter_type="box", bg_transp=True, bg_transp_refract=True, see_through=True, ground_quad=False),
    "simple_dl_bgrefract_only": lambda: scenes.simple_point_scene(160, 90, 1, filter_type="box", bg_transp_refract=True, see_through=True),
    "simple_pt_bgtransp_refract": lambda: scenes.simple_point_scene(96, 54, 2, integrator="pathtracing", filter_type="box", bg_transp=True, bg_transp_refract=True, see_through=True),
    "simple_pt_bgtransp": lambda: scenes.simple_point_scene(160, 90, 4, integrator="pathtracing", filter_type="box", bg_transp=True),
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
                       hdr_env=None, extra_lights=(), camera=None, gradient=None, beer=False, portal=False, **integ):
    """Every material and light type of the hot path in one small scene (configs 3/4 in miniature): glossy, coated_glossy,
    rough_glass, blend_mat, mirror, glass and shinydiffuse boxes on a ground quad; sun + mesh light + constant-background IBL."""
    from libyafaray_b200.interface import Interface
    y = Interface(); y.startScene(0)

    def mat(mname, typ, **kw):
        y.paramsClearAll(); y.paramsSetString("type", typ)
        for k, v in kw.items():
            if isinstance(v, bool): y.paramsSetBool(k, v)
            elif isinstance(v, int): y.paramsSetInt(k, v)
            elif isinstance(v, float): y.paramsSetFloat(k, v)
            elif isinstance(v, str): y.paramsSetString(k, v)
            else: y.paramsSetColor(k, *v)
        return y.createMaterial(mname)
    ground = mat("ground", "shinydiffusemat", color=(0.7, 0.7, 0.65))
    mats = [
        mat("gloss", "glossy", color=(0.9, 0.6, 0.2), diffuse_color=(0.5, 0.3, 0.1), diffuse_reflect=0.4, glossy_reflect=0.8, exponent=200.0, as_diffuse=False),
        mat("coated", "coated_glossy", color=(0.2, 0.5, 0.9), diffuse_color=(0.1, 0.2, 0.5), diffuse_reflect=0.5, glossy_reflect=0.7, exponent=80.0, IOR=1.8, as_diffuse=False),
        mat("rough", "rough_glass", IOR=1.5, alpha=0.3, filter_color=(0.9, 1.0, 0.9), transmit_filter=1.0,
            **({"name": "rough", "absorption": (0.9, 0.5, 0.6, 1.0), "absorption_dist": 0.8} if beer else {})),
        mat("aniso", "glossy", color=(0.8, 0.8, 0.8), diffuse_color=(0.3, 0.3, 0.3), diffuse_reflect=0.3, anisotropic=True, exp_u=20.0, exp_v=800.0, as_diffuse=True),
        mat("mirr", "shinydiffusemat", color=(0.6, 0.2, 0.2), specular_reflect=0.5, mirror_color=(1.0, 1.0, 1.0), fresnel_effect=True, IOR=1.6, diffuse_reflect=0.9),
        mat("glas", "glass", IOR=1.45, filter_color=(1.0, 0.9, 0.9), transmit_filter=0.8, fake_shadows=transp_shad,
            **({"name": "glas", "absorption": (0.4, 0.7, 0.9, 1.0), "absorption_dist": 0.5} if beer else {})),
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
    if portal:
        # a "window" of four triangles the background shines through (bgPortalLight); its own material is an ordinary diffuse one
        y.startTriMesh(31, 5, 4, False, False, 0)
        pa, pb, pc, pd, pe = (y.addVertex(-3.0, 3.2, 0.5), y.addVertex(3.0, 3.2, 0.5), y.addVertex(3.0, 3.6, 3.0), y.addVertex(-3.0, 3.6, 3.0), y.addVertex(0.0, 3.5, 1.75))
        for q in ((pa, pb, pe), (pb, pc, pe), (pc, pd, pe), (pd, pa, pe)):   # wound so that the normals face the boxes (-y)
            y.addTriangle(q[0], q[1], q[2], ground)
        y.endTriMesh()
    y.endGeometry()
    if portal:
        y.paramsClearAll(); y.paramsSetString("type", "bgPortalLight"); y.paramsSetInt("object", 31); y.paramsSetFloat("power", 2.5); y.paramsSetInt("samples", 3)
        y.createLight("portal")
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
# Adaptive AA (AA_threshold > 0, several passes): which pixels a pass resamples is ImageFilm::nextPass's decision (imagefilm.cc:306-424) — taken by the reference's own
# film in the C++ adaptor, and ON THE DEVICE (yafgpu_next_pass) in the Python mirror.  Both must reach the reference's film.
ADAPTIVE = dict(FRONTEND_ONLY)
ADAPTIVE.update({
    "cornell_dl_adaptive_dark": lambda: scenes.cornell_box(160, 90, 2, integrator="directlighting", aa_passes=4, AA_threshold=0.05, AA_inc_samples=1,
                                                           AA_dark_detection_type="linear", AA_dark_threshold_factor=0.8),
    "cornell_dl_adaptive_curve": lambda: scenes.cornell_box(160, 90, 1, integrator="directlighting", aa_passes=3, AA_threshold=0.05, AA_inc_samples=2,
                                                            AA_dark_detection_type="curve", AA_detect_color_noise=True),
    "cornell_dl_adaptive_variance": lambda: scenes.cornell_box(160, 90, 2, integrator="directlighting", aa_passes=4, AA_threshold=0.03, AA_inc_samples=1,
                                                               AA_variance_pixels=3, AA_variance_edge_size=6, AA_resampled_floor=30.0, AA_sample_multiplier_factor=1.5),
    "simple_dl_adaptive_gauss": lambda: scenes.simple_point_scene(160, 90, 1, aa_passes=3, AA_threshold=0.01, AA_inc_samples=2),
})
FRONTEND_ONLY = ADAPTIVE
CONVERGED["zoo_pt"] = (lambda spp: material_zoo_scene(W, H, spp), 8192)


# ---- §8f rank 1: shader-node graphs and image textures ---------------------------------------------------------------------
TEX_HDR = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "golden", "tex_48x40.hdr")


def make_tex_hdr(path=TEX_HDR):
    """The texture fixture (written once by tools/make_golden.py): smooth colour ramps, a diagonal stripe pattern and a few bright texels;
    non-square and not a power of two so that row/column mix-ups show."""
    import numpy as np
    from libyafaray_b200.interface import write_radiance_hdr
    h, w = 40, 48
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    img = np.zeros((h, w, 3), np.float32)
    img[..., 0] = 0.15 + 0.8 * xx / (w - 1)
    img[..., 1] = 0.1 + 0.85 * yy / (h - 1)
    img[..., 2] = 0.5 + 0.45 * np.sin(0.9 * xx + 0.6 * yy)
    stripe = ((xx + 2 * yy).astype(np.int32) // 5) % 2 == 0
    img[stripe] *= 0.35
    img[5:9, 30:36] = (1.6, 0.2, 0.1)
    write_radiance_hdr(path, img)
    return path


def textured_scene(width=96, height=54, spp=2, integrator="directlighting", tex=None, variant="basic", **integ):
    """Ground quad with uv coordinates and six boxes (three of them with orco coordinates) whose shinydiffuse materials read their
    colours / component strengths / bump from node graphs: texture_mapper (uv, orco, global, transformed; plain, cube, tube, sphere),
    layer (all blend modes, stencil, negative, noRGB), mix (all modes) and value nodes over an image texture."""
    from libyafaray_b200.interface import Interface
    y = Interface(); y.startScene(0)
    tex = dict(tex or {})
    y.paramsClearAll(); y.paramsSetString("type", "image"); y.paramsSetString("filename", TEX_HDR)
    _set_params(y, tex); y.createTexture("tex")
    y.paramsClearAll(); y.paramsSetString("type", "image"); y.paramsSetString("filename", ENV_HDR); y.paramsSetString("interpolate", "bicubic")
    y.paramsSetString("clipping", "extend"); y.createTexture("tex2")

    def node(**kw):
        y.paramsPushList(); y.paramsSetString("element", "shader_node"); _set_params(y, kw)

    def mapper(name, texture="tex", **kw):
        node(type="texture_mapper", name=name, texture=texture, **kw)

    def layer(name, inp, **kw):
        d = dict(type="layer", name=name, input=inp, mode=0, colfac=1.0, def_col=(1.0, 0.0, 1.0, 1.0), def_val=1.0, do_color=True, do_scalar=False,
                 color_input=True, negative=False, noRGB=False, stencil=False, use_alpha=False, upper_color=(0.9, 0.9, 0.9, 1.0), upper_value=0.0)
        d.update(kw)
        node(**d)

    def sd(name, **kw):
        y.paramsClearAll(); y.paramsSetString("type", "shinydiffusemat"); _set_params(y, kw)

    modes = {"basic": (0, 2, 4), "modes_b": (1, 3, 6), "modes_c": (5, 7, 8)}.get(variant, (0, 2, 4))
    if variant == "mask":
        _textured_shinydiffuse_materials(y, sd, node, mapper, layer, modes, variant)
        _textured_mask_materials(y, sd, node, mapper, layer)
        box_mats = (("mb0", True, False), ("mb1", False, False), ("b2", True, False), ("mb3", True, False), ("b4", True, False), ("b5", False, True))
    elif variant == "procedural":
        _textured_procedural_materials(y, sd, node, mapper, layer)
        box_mats = (("b0", True, False), ("b1", False, False), ("b2", True, False), ("b3", True, False), ("b4", True, False), ("b5", False, True))
    elif variant in ("noise", "noise2", "ramp"):
        {"noise": _textured_noise_materials, "noise2": _textured_noise2_materials, "ramp": _textured_ramp_materials}[variant](y, sd, node, mapper, layer)
        box_mats = (("b0", True, False), ("b1", False, False), ("b2", True, False), ("b3", True, False), ("b4", True, False), ("b5", False, True))
    elif variant == "materials":
        _textured_other_materials(y, sd, node, mapper, layer)
        box_mats = (("b0", True, False), ("b1", False, False), ("b2", True, False), ("b3", True, False), ("b4", True, False), ("b5", False, True))
    else:
        _textured_shinydiffuse_materials(y, sd, node, mapper, layer, modes, variant)
        box_mats = (("b0", True, False), ("b1", False, False), ("b2", True, False), ("b3", True, False), ("b4", True, False), ("b5", False, True))
    _textured_geometry_and_rest(y, box_mats, width, height, spp, integrator, integ)
    return y


def _textured_procedural_materials(y, sd, node, mapper, layer):
    """The procedural textures without a noise basis: "blend" (all seven progressions, flipped axis) and "rgb_cube" (with adjustments), read through
    colour layers, scalar layers and as bump maps (the three-dimensional finite-difference branch of TextureMapperNode::evalDerivative)."""
    def tex(name, typ, **kw):
        y.paramsClearAll(); y.paramsSetString("type", typ); _set_params(y, kw); y.createTexture(name)
    for st in ("lin", "quad", "ease", "diag", "sphere", "halo", "radial"):
        tex("p_" + st, "blend", stype=st)
    tex("p_flip", "blend", stype="diag", use_flip_axis=True)
    tex("p_cube", "rgb_cube")
    tex("p_cube_adj", "rgb_cube", adj_intensity=1.1, adj_contrast=1.4, adj_saturation=0.7, adj_hue=25.0, adj_mult_factor_red=0.9, adj_mult_factor_blue=1.2, adj_clamp=True)
    sd("ground", color=(0.7, 0.7, 0.7, 1.0), diffuse_shader="gl")
    layer("gl", "gm", mode=0, colfac=0.8, upper_color=(0.3, 0.5, 0.8, 1.0))
    mapper("gm", texture="p_radial", texco="uv", mapping="cube", scale=(2.0, 2.0, 1.0), offset=(-0.5, -0.5, 0.0))   # "mapping" is ignored for procedurals
    y.createMaterial("ground")
    sd("b0", color=(0.9, 0.9, 0.9, 1.0), diffuse_shader="l0")
    layer("l0", "m0", mode=0, colfac=0.9)
    mapper("m0", texture="p_cube_adj", texco="orco", scale=(0.5, 0.5, 0.5), offset=(0.25, 0.25, 0.25))
    y.createMaterial("b0")
    sd("b1", color=(0.8, 0.6, 0.4, 1.0), specular_reflect=0.3, mirror_shader="l1s", diffuse_shader="l1c")
    layer("l1s", "m1", do_color=False, do_scalar=True, color_input=False, def_val=0.6, upper_value=0.1)
    layer("l1c", "m1b", mode=2, colfac=0.7, upper_color=(0.9, 0.8, 0.3, 1.0))
    mapper("m1", texture="p_ease", texco="global", scale=(0.6, 0.6, 0.6))
    mapper("m1b", texture="p_quad", texco="global", scale=(0.4, 0.4, 0.4), offset=(0.3, 0.0, 0.0))
    y.createMaterial("b1")
    sd("b2", color=(0.5, 0.8, 0.5, 1.0), transparency=0.1, transmit_filter=0.7, transparency_shader="l2t", diffuse_shader="l2b")
    layer("l2t", "m2", do_color=False, do_scalar=True, color_input=True, def_val=0.5, upper_value=0.0)
    layer("l2b", "m2b", mode=0, colfac=0.8)
    mapper("m2", texture="p_sphere", texco="orco")
    mapper("m2b", texture="p_cube", texco="orco", scale=(0.5, 0.5, 0.5), offset=(0.25, 0.25, 0.25))
    y.createMaterial("b2")
    sd("b3", color=(0.8, 0.8, 0.8, 1.0), diffuse_shader="x3")
    node(type="mix", name="x3", input1="m3", input2="m3b", value=0.5, mode=0)
    mapper("m3", texture="p_halo", texco="orco", scale=(0.8, 0.8, 0.8))
    mapper("m3b", texture="p_flip", texco="orco")
    y.createMaterial("b3")
    sd("b4", color=(0.7, 0.7, 0.9, 1.0), specular_reflect=0.15, bump_shader="l4b", diffuse_shader="l4c")
    layer("l4b", "m4", do_color=False, do_scalar=True, color_input=False)
    layer("l4c", "m4c", mode=0, colfac=0.5, upper_color=(0.9, 0.4, 0.3, 1.0))
    mapper("m4", texture="p_sphere", texco="orco", scale=(1.5, 1.5, 1.5), bump_strength=3.0)
    mapper("m4c", texture="p_lin", texco="orco")
    y.createMaterial("b4")
    sd("b5", color=(0.9, 0.8, 0.6, 1.0), bump_shader="l5b", diffuse_shader="l5c")
    layer("l5b", "m5", do_color=False, do_scalar=True, color_input=False)
    layer("l5c", "m5c", mode=0, colfac=0.8)
    mapper("m5", texture="p_halo", texco="uv", scale=(2.0, 2.0, 1.0), offset=(-0.5, -0.5, 0.0), bump_strength=2.0)   # uv coordinates, yet the non-discrete bump branch
    mapper("m5c", texture="p_diag", texco="uv")
    y.createMaterial("b5")


def _textured_noise_materials(y, sd, node, mapper, layer):
    """The noise-based procedural textures: "clouds", "marble", "wood" (texture_basic.cc:96-343) over every basis newNoise__ can build (new Perlin,
    standard Perlin, Blender noise, six Voronoi variants, cell noise), soft and hard turbulence, the three wave shapes, ring wood, cloud bias,
    adjustments; read as colours, as scalars and as bump maps."""
    def tex(name, typ, **kw):
        y.paramsClearAll(); y.paramsSetString("type", typ); _set_params(y, kw); y.createTexture(name)
    tex("n_clouds_new", "clouds", noise_type="newperlin", depth=3, size=2.5, color1=(0.1, 0.2, 0.6, 1.0), color2=(1.0, 0.9, 0.7, 1.0))
    tex("n_clouds_std", "clouds", noise_type="stdperlin", depth=2, size=4.0, hard=True)
    tex("n_clouds_bl", "clouds", noise_type="blender", depth=4, size=3.0, bias="negative", adj_intensity=1.2, adj_contrast=1.5, adj_clamp=True)
    tex("n_clouds_pos", "clouds", noise_type="blender", depth=1, size=2.0, bias="positive")
    tex("n_clouds_cell", "clouds", noise_type="cellnoise", depth=0, size=6.0, color1=(0.9, 0.1, 0.1, 1.0), color2=(0.1, 0.9, 0.3, 1.0))
    tex("n_marble_sin", "marble", noise_type="blender", depth=3, size=2.0, turbulence=4.0, sharpness=3.0, color1=(0.9, 0.9, 0.85, 1.0), color2=(0.2, 0.25, 0.3, 1.0))
    tex("n_marble_saw", "marble", noise_type="voronoi_f1", depth=1, size=3.0, turbulence=2.0, shape="saw", hard=True)
    tex("n_marble_tri", "marble", noise_type="voronoi_crackle", depth=2, size=2.0, turbulence=3.0, shape="tri", sharpness=0.5, adj_saturation=0.6, adj_hue=40.0)
    tex("n_marble_flat", "marble", noise_type="newperlin", turbulence=0.0, size=1.0, color1=(0.8, 0.5, 0.2, 1.0), color2=(0.1, 0.1, 0.4, 1.0))
    tex("n_wood_bands", "wood", noise_type="stdperlin", depth=2, size=3.0, turbulence=2.0, color1=(0.45, 0.25, 0.1, 1.0), color2=(0.8, 0.6, 0.35, 1.0),
        adj_intensity=1.3, adj_contrast=1.2)   # read as a colour: intensity / contrast act on the scalar only, the colour gets applyColorAdjustments
    tex("n_wood_rings", "wood", noise_type="voronoi_f2f1", depth=1, size=2.0, turbulence=1.5, wood_type="rings", shape="tri")
    tex("n_wood_saw", "wood", noise_type="voronoi_f2", depth=0, size=1.5, turbulence=0.5, shape="saw", adj_mult_factor_red=1.2, adj_mult_factor_green=0.8)
    tex("n_clouds_f3", "clouds", noise_type="voronoi_f3", depth=1, size=3.0)
    tex("n_clouds_f4", "clouds", noise_type="voronoi_f4", depth=2, size=2.0, hard=True)
    sd("ground", color=(0.7, 0.7, 0.7, 1.0), diffuse_shader="gl")
    layer("gl", "gm", mode=0, colfac=0.9, upper_color=(0.3, 0.5, 0.8, 1.0))
    mapper("gm", texture="n_marble_sin", texco="global", scale=(0.3, 0.3, 0.3))
    y.createMaterial("ground")
    sd("b0", color=(0.9, 0.9, 0.9, 1.0), diffuse_shader="l0")
    layer("l0", "m0", mode=0, colfac=0.9)
    mapper("m0", texture="n_clouds_new", texco="orco", scale=(1.5, 1.5, 1.5), offset=(0.25, 0.25, 0.25))
    y.createMaterial("b0")
    sd("b1", color=(0.8, 0.6, 0.4, 1.0), specular_reflect=0.3, mirror_shader="l1s", diffuse_shader="l1c")
    layer("l1s", "m1", do_color=False, do_scalar=True, color_input=False, def_val=0.6, upper_value=0.1)
    layer("l1c", "m1b", mode=2, colfac=0.7, upper_color=(0.9, 0.8, 0.3, 1.0))
    mapper("m1", texture="n_clouds_std", texco="global", scale=(0.6, 0.6, 0.6))
    mapper("m1b", texture="n_wood_bands", texco="global", scale=(0.4, 0.4, 0.4), offset=(0.3, 0.0, 0.0))
    y.createMaterial("b1")
    sd("b2", color=(0.5, 0.8, 0.5, 1.0), transparency=0.1, transmit_filter=0.7, transparency_shader="l2t", diffuse_shader="l2b")
    layer("l2t", "m2", do_color=False, do_scalar=True, color_input=True, def_val=0.5, upper_value=0.0)
    layer("l2b", "m2b", mode=0, colfac=0.8)
    mapper("m2", texture="n_clouds_pos", texco="orco", scale=(2.0, 2.0, 2.0))
    mapper("m2b", texture="n_marble_tri", texco="orco", scale=(1.5, 1.5, 1.5), offset=(0.25, 0.25, 0.25))
    y.createMaterial("b2")
    sd("b3", color=(0.8, 0.8, 0.8, 1.0), diffuse_shader="x3")
    node(type="mix", name="x3", input1="m3", input2="m3b", value=0.5, mode=0)
    mapper("m3", texture="n_marble_saw", texco="orco", scale=(1.8, 1.8, 1.8))
    mapper("m3b", texture="n_clouds_cell", texco="orco")
    y.createMaterial("b3")
    sd("b4", color=(0.7, 0.7, 0.9, 1.0), specular_reflect=0.15, bump_shader="l4b", diffuse_shader="x4")
    layer("l4b", "m4", do_color=False, do_scalar=True, color_input=False)
    node(type="mix", name="x4", input1="m4c", input2="m4d", value=0.4, mode=1)
    mapper("m4", texture="n_clouds_bl", texco="orco", scale=(1.5, 1.5, 1.5), bump_strength=3.0)
    mapper("m4c", texture="n_wood_rings", texco="orco")
    mapper("m4d", texture="n_clouds_f3", texco="orco", scale=(2.0, 2.0, 2.0))
    y.createMaterial("b4")
    sd("b5", color=(0.9, 0.8, 0.6, 1.0), bump_shader="l5b", diffuse_shader="x5")
    layer("l5b", "m5", do_color=False, do_scalar=True, color_input=False)
    node(type="mix", name="x5", input1="m5c", input2="m5d", value=0.5, mode=0)
    mapper("m5", texture="n_marble_flat", texco="uv", scale=(2.0, 2.0, 1.0), offset=(-0.5, -0.5, 0.0), bump_strength=2.0)
    mapper("m5c", texture="n_wood_saw", texco="uv", scale=(3.0, 3.0, 1.0))
    mapper("m5d", texture="n_clouds_f4", texco="uv", scale=(4.0, 4.0, 1.0))
    y.createMaterial("b5")


def _textured_noise2_materials(y, sd, node, mapper, layer):
    """The other noise-based procedurals: "voronoi" (all distance metrics, intensity and the three cell-colour modes), "musgrave" (five fractals, fractional
    octaves) and "distorted_noise" (texture_basic.cc:391-673)."""
    def tex(name, typ, **kw):
        y.paramsClearAll(); y.paramsSetString("type", typ); _set_params(y, kw); y.createTexture(name)
    tex("v_int", "voronoi", size=3.0, weight1=1.0, weight2=-0.5, intensity=1.5)
    tex("v_col1", "voronoi", size=2.5, Rgbype="col1", weight1=0.6, weight2=0.3, weight3=0.2, weight4=0.1, distance_metric="squared")
    tex("v_col2", "voronoi", size=2.0, Rgbype="col2", weight1=1.0, weight2=0.5, distance_metric="manhattan", adj_intensity=1.1, adj_contrast=1.3, adj_saturation=0.8)
    tex("v_col3", "voronoi", size=3.5, Rgbype="col3", weight1=0.2, weight2=1.0, distance_metric="chebychev", intensity=2.0)
    tex("v_mh", "voronoi", size=2.0, distance_metric="minkovsky_half", weight1=0.7, weight3=0.4)
    tex("v_m4", "voronoi", size=2.0, distance_metric="minkovsky_four", weight2=1.0, weight1=0.0)
    tex("v_mk", "voronoi", size=1.5, distance_metric="minkovsky", mk_exponent=3.5, weight1=1.0, weight4=0.5)
    tex("m_fbm", "musgrave", noise_type="newperlin", musgrave_type="fBm", H=0.8, lacunarity=2.1, octaves=3.5, size=2.0, intensity=0.8,
        color1=(0.1, 0.1, 0.5, 1.0), color2=(1.0, 0.8, 0.4, 1.0))
    tex("m_mf", "musgrave", noise_type="blender", musgrave_type="multifractal", H=1.0, lacunarity=2.0, octaves=4.0, size=1.5)
    tex("m_ht", "musgrave", noise_type="stdperlin", musgrave_type="heteroterrain", H=0.9, lacunarity=2.3, octaves=2.25, offset=0.7, size=2.0, intensity=0.5)
    tex("m_hy", "musgrave", noise_type="voronoi_f1", musgrave_type="hybridmf", H=0.25, lacunarity=2.0, octaves=5.5, offset=0.7, gain=1.2, size=1.2, adj_hue=30.0)
    tex("m_ri", "musgrave", noise_type="cellnoise", musgrave_type="ridgedmf", H=1.0, lacunarity=2.0, octaves=4.0, offset=1.0, gain=2.0, size=3.0, intensity=0.6)
    tex("d_a", "distorted_noise", noise_type1="blender", noise_type2="newperlin", distort=1.5, size=2.0, color1=(0.9, 0.3, 0.1, 1.0), color2=(0.1, 0.3, 0.9, 1.0))
    tex("d_b", "distorted_noise", noise_type1="voronoi_f2f1", noise_type2="stdperlin", distort=0.7, size=3.0, adj_mult_factor_blue=1.3, adj_clamp=True)
    sd("ground", color=(0.7, 0.7, 0.7, 1.0), diffuse_shader="gl")
    layer("gl", "gm", mode=0, colfac=0.9, upper_color=(0.3, 0.5, 0.8, 1.0))
    mapper("gm", texture="v_col1", texco="global", scale=(0.3, 0.3, 0.3))
    y.createMaterial("ground")
    sd("b0", color=(0.9, 0.9, 0.9, 1.0), diffuse_shader="x0")
    node(type="mix", name="x0", input1="m0", input2="m0b", value=0.5, mode=0)
    mapper("m0", texture="m_fbm", texco="orco", scale=(1.5, 1.5, 1.5), offset=(0.25, 0.25, 0.25))
    mapper("m0b", texture="v_mk", texco="orco")
    y.createMaterial("b0")
    sd("b1", color=(0.8, 0.6, 0.4, 1.0), specular_reflect=0.3, mirror_shader="l1s", diffuse_shader="x1")
    layer("l1s", "m1", do_color=False, do_scalar=True, color_input=False, def_val=0.6, upper_value=0.1)
    node(type="mix", name="x1", input1="m1b", input2="m1c", value=0.6, mode=0)
    mapper("m1", texture="v_int", texco="global", scale=(0.6, 0.6, 0.6))
    mapper("m1b", texture="m_mf", texco="global", scale=(0.4, 0.4, 0.4), offset=(0.3, 0.0, 0.0))
    mapper("m1c", texture="v_col3", texco="global", scale=(0.5, 0.5, 0.5))
    y.createMaterial("b1")
    sd("b2", color=(0.5, 0.8, 0.5, 1.0), transparency=0.1, transmit_filter=0.7, transparency_shader="l2t", diffuse_shader="x2")
    layer("l2t", "m2", do_color=False, do_scalar=True, color_input=True, def_val=0.5, upper_value=0.0)
    node(type="mix", name="x2", input1="m2b", input2="m2c", value=0.5, mode=0)
    mapper("m2", texture="m_ht", texco="orco", scale=(2.0, 2.0, 2.0))
    mapper("m2b", texture="d_a", texco="orco", scale=(1.5, 1.5, 1.5), offset=(0.25, 0.25, 0.25))
    mapper("m2c", texture="v_mh", texco="orco")
    y.createMaterial("b2")
    sd("b3", color=(0.8, 0.8, 0.8, 1.0), diffuse_shader="x3")
    node(type="mix", name="x3", input1="m3", input2="m3b", value=0.5, mode=0)
    mapper("m3", texture="v_col2", texco="orco", scale=(1.8, 1.8, 1.8))
    mapper("m3b", texture="m_hy", texco="orco")
    y.createMaterial("b3")
    sd("b4", color=(0.7, 0.7, 0.9, 1.0), specular_reflect=0.15, bump_shader="l4b", diffuse_shader="x4")
    layer("l4b", "m4", do_color=False, do_scalar=True, color_input=False)
    node(type="mix", name="x4", input1="m4c", input2="m4d", value=0.4, mode=1)
    mapper("m4", texture="m_ri", texco="orco", scale=(1.5, 1.5, 1.5), bump_strength=3.0)
    mapper("m4c", texture="d_b", texco="orco")
    mapper("m4d", texture="v_m4", texco="orco", scale=(2.0, 2.0, 2.0))
    y.createMaterial("b4")
    sd("b5", color=(0.9, 0.8, 0.6, 1.0), bump_shader="l5b", diffuse_shader="l5c")
    layer("l5b", "m5", do_color=False, do_scalar=True, color_input=False)
    layer("l5c", "m5c", mode=0, colfac=0.8, use_alpha=True)
    mapper("m5", texture="v_int", texco="uv", scale=(2.0, 2.0, 1.0), offset=(-0.5, -0.5, 0.0), bump_strength=2.0)
    mapper("m5c", texture="v_col1", texco="uv", scale=(3.0, 3.0, 1.0))
    y.createMaterial("b5")


def _textured_ramp_materials(y, sd, node, mapper, layer):
    """Colour ramps (ColorRamp::getColorInterpolated, color_ramp.cc:84-164) on every texture type that reads one: RGB linear / constant, HSV and HSL with the four hue
    interpolations, ramp alpha through a use_alpha layer, items given out of order, scalars beyond [0, 1] (negated cloud bias, Musgrave)."""
    def tex(name, typ, ramp, mode="RGB", interp="LINEAR", hue="NEAR", **kw):
        y.paramsClearAll(); y.paramsSetString("type", typ); _set_params(y, kw)
        y.paramsSetBool("use_color_ramp", True); y.paramsSetString("ramp_color_mode", mode); y.paramsSetString("ramp_interpolation", interp)
        y.paramsSetString("ramp_hue_interpolation", hue); y.paramsSetInt("ramp_num_items", len(ramp))
        for i, (pos, col) in enumerate(ramp):
            y.paramsSetColor("ramp_item_%d_color" % i, col[0], col[1], col[2], 1.0); y.paramsSetFloat("ramp_item_%d_alpha" % i, col[3]); y.paramsSetFloat("ramp_item_%d_position" % i, pos)
        y.createTexture(name)
    fire = [(0.0, (0.05, 0.0, 0.0, 1.0)), (0.4, (0.9, 0.1, 0.0, 1.0)), (0.7, (1.0, 0.8, 0.1, 0.6)), (1.0, (1.0, 1.0, 0.9, 0.2))]
    unordered = [(1.0, (0.1, 0.2, 0.9, 1.0)), (0.0, (0.9, 0.9, 0.2, 1.0)), (0.5, (0.1, 0.8, 0.3, 0.5))]
    two = [(0.0, (0.9, 0.1, 0.1, 1.0)), (1.0, (0.1, 0.1, 0.9, 1.0))]
    tex("r_blend", "blend", fire, stype="lin")
    tex("r_clouds", "clouds", unordered, noise_type="blender", depth=2, size=3.0)
    tex("r_clouds_neg", "clouds", fire, interp="CONSTANT", noise_type="newperlin", depth=1, size=2.0, bias="positive")
    tex("r_marble", "marble", two, mode="HSV", hue="NEAR", noise_type="stdperlin", depth=2, size=2.0, turbulence=3.0)
    tex("r_wood", "wood", two, mode="HSV", hue="FAR", noise_type="blender", depth=1, size=2.0, turbulence=1.0, adj_saturation=0.8)
    tex("r_voronoi", "voronoi", unordered, mode="HSV", hue="CW", size=3.0, weight1=1.0, weight2=0.3)
    tex("r_musgrave", "musgrave", fire, mode="HSV", hue="CCW", noise_type="newperlin", musgrave_type="fBm", octaves=3.0, size=2.0, intensity=0.7)
    tex("r_dist", "distorted_noise", unordered, mode="HSL", noise_type1="blender", noise_type2="stdperlin", distort=1.0, size=2.0)
    tex("r_blend_hsl", "blend", two, mode="hsl", hue="ccw", stype="sphere")
    sd("ground", color=(0.7, 0.7, 0.7, 1.0), diffuse_shader="gl")
    layer("gl", "gm", mode=0, colfac=0.9, upper_color=(0.3, 0.5, 0.8, 1.0))
    mapper("gm", texture="r_marble", texco="global", scale=(0.3, 0.3, 0.3))
    y.createMaterial("ground")
    sd("b0", color=(0.9, 0.9, 0.9, 1.0), diffuse_shader="l0")
    layer("l0", "m0", mode=0, colfac=0.9, use_alpha=True, upper_color=(0.2, 0.2, 0.2, 1.0))
    mapper("m0", texture="r_blend", texco="orco")
    y.createMaterial("b0")
    sd("b1", color=(0.8, 0.6, 0.4, 1.0), diffuse_shader="x1")
    node(type="mix", name="x1", input1="m1", input2="m1b", value=0.5, mode=0)
    mapper("m1", texture="r_clouds", texco="global", scale=(0.6, 0.6, 0.6))
    mapper("m1b", texture="r_clouds_neg", texco="global", scale=(0.4, 0.4, 0.4), offset=(0.3, 0.0, 0.0))
    y.createMaterial("b1")
    sd("b2", color=(0.5, 0.8, 0.5, 1.0), diffuse_shader="l2b")
    layer("l2b", "m2b", mode=0, colfac=0.8)
    mapper("m2b", texture="r_wood", texco="orco", scale=(1.5, 1.5, 1.5), offset=(0.25, 0.25, 0.25))
    y.createMaterial("b2")
    sd("b3", color=(0.8, 0.8, 0.8, 1.0), diffuse_shader="x3")
    node(type="mix", name="x3", input1="m3", input2="m3b", value=0.5, mode=0)
    mapper("m3", texture="r_voronoi", texco="orco", scale=(1.8, 1.8, 1.8))
    mapper("m3b", texture="r_musgrave", texco="orco")
    y.createMaterial("b3")
    sd("b4", color=(0.7, 0.7, 0.9, 1.0), diffuse_shader="l4c")
    layer("l4c", "m4c", mode=0, colfac=0.9, use_alpha=True, upper_color=(0.9, 0.4, 0.3, 1.0))
    mapper("m4c", texture="r_dist", texco="orco")
    y.createMaterial("b4")
    sd("b5", color=(0.9, 0.8, 0.6, 1.0), diffuse_shader="l5c")
    layer("l5c", "m5c", mode=0, colfac=0.8)
    mapper("m5c", texture="r_blend_hsl", texco="uv", scale=(2.0, 2.0, 1.0), offset=(-0.5, -0.5, 0.0))
    y.createMaterial("b5")


def _textured_mask_materials(y, sd, node, mapper, layer):
    """mask_mat (material_mask.cc): a textured scalar against a threshold picks one of two materials per hit; used on boxes 0, 1 and 3."""
    # MaskMaterial::initBsdf evaluates all_nodes_ in DECLARATION order (material_mask.cc:43), not in dependency order: a node declared before its
    # input reads whatever the previous hit left in the stack.  The nodes are therefore declared inputs first (the device always uses dependency order).
    def mat(typ, **kw):
        y.paramsClearAll(); y.paramsSetString("type", typ); _set_params(y, kw)
    mat("glossy", color=(0.9, 0.8, 0.3, 1.0), diffuse_color=(0.4, 0.3, 0.1, 1.0), diffuse_reflect=0.4, glossy_reflect=0.8, exponent=200.0, as_diffuse=False)
    y.createMaterial("mk_a")
    mat("shinydiffusemat", color=(0.2, 0.3, 0.8, 1.0), specular_reflect=0.3, diffuse_reflect=0.9)
    y.createMaterial("mk_b")
    mat("glass", IOR=1.5, filter_color=(0.9, 1.0, 0.9, 1.0), transmit_filter=0.9)
    y.createMaterial("mk_g")
    mat("shinydiffusemat", color=(0.8, 0.25, 0.2, 1.0), diffuse_brdf="oren_nayar", sigma=0.3)
    y.createMaterial("mk_l")
    mat("blend_mat", material1="mk_a", material2="mk_l", blend_value=0.4)
    y.createMaterial("mk_bl")
    mat("mask_mat", material1="mk_a", material2="mk_b", threshold=0.45, mask="k0_l")
    mapper("k0_m", texco="orco", mapping="cube", scale=(1.2, 1.2, 1.2))
    layer("k0_l", "k0_m", do_color=False, do_scalar=True, color_input=True, def_val=1.0, upper_value=0.0)
    y.createMaterial("mb0")
    mat("mask_mat", material1="mk_g", material2="mk_bl", threshold=0.35, mask="k1_l", additionaldepth=1)
    mapper("k1_m", texco="global", mapping="tube")
    layer("k1_l", "k1_m", do_color=False, do_scalar=True, color_input=False, def_val=1.0, upper_value=0.0)
    y.createMaterial("mb1")
    mat("mask_mat", material1="mk_l", material2="mk_a", threshold=0.5, mask="k3_l", receive_shadows=False)
    mapper("k3_m", texco="orco", mapping="sphere")
    layer("k3_l", "k3_m", do_color=False, do_scalar=True, color_input=True, def_val=1.0, upper_value=0.0, negative=True)
    y.createMaterial("mb3")


def _textured_other_materials(y, sd, node, mapper, layer):
    """glossy, coated_glossy, glass, rough_glass and blend_mat with every "<slot>_shader" those factories know."""
    def mat(typ, **kw):
        y.paramsClearAll(); y.paramsSetString("type", typ); _set_params(y, kw)
    sd("ground", color=(0.7, 0.7, 0.7, 1.0), diffuse_shader="gl")
    layer("gl", "gm", mode=0, colfac=0.9)
    mapper("gm", texco="uv", mapping="plain", scale=(3.0, 2.0, 1.0))
    y.createMaterial("ground")
    # glossy: glossy / diffuse colours, reflectivity, exponent, diffuse_refl and a textured Oren-Nayar sigma
    mat("glossy", color=(0.9, 0.6, 0.2, 1.0), diffuse_color=(0.5, 0.3, 0.1, 1.0), diffuse_reflect=0.5, glossy_reflect=0.7, exponent=120.0, as_diffuse=False,
        diffuse_brdf="Oren-Nayar", sigma=0.2, glossy_shader="g_c", diffuse_shader="g_d", glossy_reflect_shader="g_r", exponent_shader="g_e",
        diffuse_refl_shader="g_f", sigma_oren_shader="g_s")
    layer("g_c", "g_m", mode=0, colfac=0.6, upper_color=(0.9, 0.6, 0.2, 1.0))
    layer("g_d", "g_m", mode=2, colfac=0.8, upper_color=(0.5, 0.4, 0.3, 1.0))
    layer("g_r", "g_m", do_color=False, do_scalar=True, color_input=True, def_val=0.9, upper_value=0.3)
    layer("g_e", "g_m", do_color=False, do_scalar=True, color_input=True, def_val=300.0, upper_value=20.0)
    layer("g_f", "g_m", do_color=False, do_scalar=True, color_input=False, def_val=0.9, upper_value=0.4)
    layer("g_s", "g_m", do_color=False, do_scalar=True, color_input=True, def_val=0.7, upper_value=0.05)
    mapper("g_m", texco="orco", mapping="cube", scale=(1.3, 1.3, 1.3))
    y.createMaterial("b0")
    # coated glossy: + mirror strength / colour and IOR from nodes, bump
    mat("coated_glossy", color=(0.2, 0.5, 0.9, 1.0), diffuse_color=(0.1, 0.2, 0.5, 1.0), diffuse_reflect=0.5, glossy_reflect=0.6, exponent=60.0, IOR=1.6, as_diffuse=False,
        mirror_color=(1.0, 1.0, 1.0, 1.0), specular_reflect=0.8, glossy_shader="c_c", diffuse_shader="c_d", mirror_shader="c_ms", mirror_color_shader="c_mc",
        IOR_shader="c_i", bump_shader="c_b", exponent_shader="c_e")
    layer("c_c", "c_m", mode=0, colfac=0.5, upper_color=(0.2, 0.5, 0.9, 1.0))
    layer("c_d", "c_m", mode=4, colfac=0.7, upper_color=(0.3, 0.3, 0.6, 1.0))
    layer("c_ms", "c_m", do_color=False, do_scalar=True, color_input=True, def_val=1.0, upper_value=0.3)
    layer("c_mc", "c_m", mode=0, colfac=0.4, upper_color=(1.0, 0.9, 0.8, 1.0))
    layer("c_i", "c_m", do_color=False, do_scalar=True, color_input=True, def_val=0.5, upper_value=0.0)
    layer("c_b", "c_m", do_color=False, do_scalar=True, color_input=False)
    layer("c_e", "c_m", do_color=False, do_scalar=True, color_input=False, def_val=200.0, upper_value=30.0)
    mapper("c_m", texco="global", mapping="plain", scale=(0.9, 0.9, 0.9), bump_strength=2.0)
    y.createMaterial("b1")
    # glass: filter and mirror colours, IOR, bump
    mat("glass", IOR=1.45, filter_color=(1.0, 0.9, 0.9, 1.0), transmit_filter=0.8, mirror_color=(1.0, 1.0, 1.0, 1.0), filter_color_shader="q_f", mirror_color_shader="q_m",
        IOR_shader="q_i", bump_shader="q_b")
    layer("q_f", "q_t", mode=0, colfac=0.5, upper_color=(0.95, 0.95, 0.95, 1.0))
    layer("q_m", "q_t", mode=0, colfac=0.3, upper_color=(1.0, 1.0, 1.0, 1.0))
    layer("q_i", "q_t", do_color=False, do_scalar=True, color_input=True, def_val=0.3, upper_value=0.0)
    layer("q_b", "q_t", do_color=False, do_scalar=True, color_input=False)
    mapper("q_t", texco="orco", mapping="cube", bump_strength=1.5)
    y.createMaterial("b2")
    # rough glass: + roughness
    mat("rough_glass", IOR=1.5, alpha=0.3, filter_color=(0.9, 1.0, 0.9, 1.0), transmit_filter=1.0, filter_color_shader="r_f", mirror_color_shader="r_m", roughness_shader="r_r",
        IOR_shader="r_i")
    layer("r_f", "r_t", mode=0, colfac=0.4, upper_color=(0.9, 1.0, 0.9, 1.0))
    layer("r_m", "r_t", mode=0, colfac=0.2, upper_color=(1.0, 1.0, 1.0, 1.0))
    layer("r_r", "r_t", do_color=False, do_scalar=True, color_input=True, def_val=0.4, upper_value=0.05)
    layer("r_i", "r_t", do_color=False, do_scalar=True, color_input=True, def_val=0.2, upper_value=0.0)
    mapper("r_t", texco="orco", mapping="sphere")
    y.createMaterial("b3")
    # blend_mat with a textured mask between two plain materials
    mat("glossy", color=(0.9, 0.8, 0.3, 1.0), diffuse_color=(0.4, 0.3, 0.1, 1.0), diffuse_reflect=0.4, glossy_reflect=0.8, exponent=200.0, as_diffuse=False)
    y.createMaterial("bl_a")
    mat("shinydiffusemat", color=(0.2, 0.3, 0.8, 1.0), specular_reflect=0.3, diffuse_reflect=0.9)
    y.createMaterial("bl_b")
    mat("blend_mat", material1="bl_a", material2="bl_b", blend_value=0.5, mask="k_l")
    layer("k_l", "k_m", do_color=False, do_scalar=True, color_input=True, def_val=1.0, upper_value=0.0)
    mapper("k_m", texco="orco", mapping="cube", scale=(1.2, 1.2, 1.2))
    y.createMaterial("b4")
    # glossy, anisotropic, as_diffuse, uv bump
    mat("glossy", color=(0.8, 0.8, 0.8, 1.0), diffuse_color=(0.3, 0.3, 0.3, 1.0), diffuse_reflect=0.3, anisotropic=True, exp_u=20.0, exp_v=800.0, as_diffuse=True,
        bump_shader="a_b", diffuse_shader="a_d")
    layer("a_b", "a_m", do_color=False, do_scalar=True, color_input=False)
    layer("a_d", "a_m", mode=0, colfac=0.9)
    mapper("a_m", texco="uv", mapping="plain", bump_strength=1.0)
    y.createMaterial("b5")


def _textured_shinydiffuse_materials(y, sd, node, mapper, layer, modes, variant):
    # ground: uv-mapped, repeated 3x2, plain projection
    sd("ground", color=(0.7, 0.7, 0.7, 1.0), diffuse_shader="gl")
    layer("gl", "gm", mode=modes[0], colfac=0.9)
    mapper("gm", texco="uv", mapping="plain", scale=(3.0, 2.0, 1.0), offset=(0.1, 0.05, 0.0))
    y.createMaterial("ground")
    # box 0: orco + cube projection (the test01 setup), with emission from the textured colour
    sd("b0", color=(0.9, 0.9, 0.9, 1.0), diffuse_reflect=0.8, emit=0.05, diffuse_shader="l0")
    layer("l0", "m0", mode=modes[1], colfac=0.8)
    mapper("m0", texco="orco", mapping="cube", proj_x=1, proj_y=2, proj_z=3)
    y.createMaterial("b0")
    # box 1: global coordinates + tube projection; mirror strength and mirror colour from nodes; Fresnel with a textured IOR
    sd("b1", color=(0.8, 0.6, 0.4, 1.0), specular_reflect=0.2, fresnel_effect=True, IOR=1.5, mirror_shader="l1s", mirror_color_shader="l1c", diffuse_shader="l1c",
       IOR_shader="l1i")
    layer("l1s", "m1", do_color=False, do_scalar=True, color_input=False, def_val=0.6, upper_value=0.1, mode=modes[2], valfac=0.9)
    layer("l1c", "m1", mode=0, colfac=0.7, upper_color=(0.3, 0.5, 0.9, 1.0))
    layer("l1i", "m1", do_color=False, do_scalar=True, color_input=True, def_val=0.4, upper_value=0.05, noRGB=False)
    mapper("m1", texco="global", mapping="tube", scale=(1.0, 1.0, 0.7))
    y.createMaterial("b1")
    # box 2: transformed... kept as global + sphere projection; transparency and translucency from nodes, stencil + negative layers stacked
    sd("b2", color=(0.5, 0.8, 0.5, 1.0), transparency=0.1, translucency=0.1, transmit_filter=0.7, transparency_shader="l2t", translucency_shader="l2u", diffuse_shader="l2b")
    layer("l2t", "m2", do_color=False, do_scalar=True, color_input=True, noRGB=True, def_val=0.5, upper_value=0.0)
    layer("l2u", "m2", do_color=False, do_scalar=True, color_input=False, def_val=0.3, upper_value=0.05, negative=True)
    layer("l2a", "m2", mode=0, stencil=True, colfac=1.0, upper_color=(0.9, 0.2, 0.2, 1.0))
    layer("l2b", "m2b", upper_layer="l2a", mode=modes[0], colfac=0.6, negative=True)
    mapper("m2", texco="global", mapping="sphere", proj_x=2, proj_y=3, proj_z=1)
    mapper("m2b", texture="tex2", texco="orco", mapping="plain", scale=(0.8, 0.8, 0.8))
    y.createMaterial("b2")
    # box 3: mix / value nodes: diffuse = mix_k(texture, value); Oren-Nayar with a textured sigma; diffuse_refl_shader
    sd("b3", color=(0.8, 0.8, 0.8, 1.0), diffuse_brdf="oren_nayar", sigma=0.3, diffuse_shader="x3", sigma_oren_shader="l3s", diffuse_refl_shader="l3r")
    node(type="mix", name="x3", input1="m3", input2="v3", factor="l3s", mode={"basic": 0, "modes_b": 4, "modes_c": 8}.get(variant, 0))
    node(type="mix", name="x3b", input1="m3", color2=(0.2, 0.3, 0.9, 1.0), value=0.4, mode={"basic": 1, "modes_b": 6, "modes_c": 7}.get(variant, 1))
    node(type="value", name="v3", color=(0.9, 0.7, 0.2, 1.0), alpha=0.8, scalar=0.35)
    layer("l3s", "m3", do_color=False, do_scalar=True, color_input=True, def_val=0.8, upper_value=0.1)
    layer("l3r", "x3b", do_color=False, do_scalar=True, color_input=True, def_val=0.9, upper_value=0.2)
    mapper("m3", texco="orco", mapping="cube", proj_x=3, proj_y=1, proj_z=2, scale=(1.5, 1.5, 1.5))
    y.createMaterial("b3")
    # box 4: bump mapping from the texture through orco coordinates (the non-uv derivative path) + a coloured layer
    sd("b4", color=(0.7, 0.7, 0.9, 1.0), specular_reflect=0.15, bump_shader="l4b", diffuse_shader="l4c")
    layer("l4b", "m4", do_color=False, do_scalar=True, color_input=False)
    layer("l4c", "m4", mode=2, colfac=0.5)
    mapper("m4", texco="orco", mapping="cube", bump_strength=3.0, scale=(2.0, 2.0, 2.0))
    y.createMaterial("b4")
    if variant == "camera_coords":
        # camera-dependent texture coordinates on two of the boxes: "window" (Camera::screenproject) and "normal" (Camera::getAxis)
        sd("b3", color=(0.8, 0.8, 0.8, 1.0), diffuse_shader="w_l")
        layer("w_l", "w_m", mode=0, colfac=0.9)
        mapper("w_m", texco="window", mapping="plain", scale=(1.5, 1.5, 1.0))
        y.createMaterial("b3")
        sd("b4", color=(0.7, 0.7, 0.9, 1.0), diffuse_shader="n_l", bump_shader="n_b")
        layer("n_l", "n_m", mode=0, colfac=0.9)
        layer("n_b", "n_m2", do_color=False, do_scalar=True, color_input=False)
        mapper("n_m", texco="normal", mapping="plain")
        mapper("n_m2", texco="orco", mapping="cube", bump_strength=2.0)
        y.createMaterial("b4")
    # box 5 (has uv): bump through uv coordinates (the uv derivative path), two stacked bump layers
    sd("b5", color=(0.9, 0.8, 0.6, 1.0), bump_shader="l5b2", diffuse_shader="l5c")
    layer("l5b1", "m5", do_color=False, do_scalar=True, color_input=False)
    layer("l5b2", "m5b", upper_layer="l5b1", do_color=False, do_scalar=True, color_input=False, negative=True)
    layer("l5c", "m5", mode=0, colfac=0.85)
    mapper("m5", texco="uv", mapping="plain", bump_strength=2.0)
    mapper("m5b", texture="tex2", texco="uv", mapping="plain", bump_strength=0.5, scale=(2.0, 2.0, 1.0))
    y.createMaterial("b5")



def _textured_geometry_and_rest(y, box_mats, width, height, spp, integrator, integ):
    def box(oid, lo, hi, mat, orco=False, uv=False, rot=0.0):
        import math
        cx, cy = 0.5 * (lo[0] + hi[0]), 0.5 * (lo[1] + hi[1])
        c, s = math.cos(math.radians(rot)), math.sin(math.radians(rot))
        y.startTriMesh(oid, 8, 12, orco, uv, 0)
        vs = []
        for k in range(8):
            px, py, pz = (hi if k & 1 else lo)[0], (hi if k & 2 else lo)[1], (hi if k & 4 else lo)[2]
            rx, ry = cx + c * (px - cx) - s * (py - cy), cy + s * (px - cx) + c * (py - cy)
            if orco:
                vs.append(y.addVertex(rx, ry, pz, 2.0 * (k & 1) - 1.0, 2.0 * ((k >> 1) & 1) - 1.0, 2.0 * ((k >> 2) & 1) - 1.0))
            else:
                vs.append(y.addVertex(rx, ry, pz))
        if uv:
            uvs = [y.addUv(0.0, 0.0), y.addUv(1.0, 0.0), y.addUv(1.0, 1.0), y.addUv(0.0, 1.0)]
        for q in ((0, 2, 3, 1), (4, 5, 7, 6), (0, 1, 5, 4), (2, 6, 7, 3), (0, 4, 6, 2), (1, 3, 7, 5)):
            a, b, cc, d = (vs[i] for i in q)
            if uv:
                y.addTriangle(a, b, cc, uvs[0], uvs[1], uvs[2], mat); y.addTriangle(a, cc, d, uvs[0], uvs[2], uvs[3], mat)
            else:
                y.addTriangle(a, b, cc, mat); y.addTriangle(a, cc, d, mat)
        y.endTriMesh()

    y.startGeometry()
    y.startTriMesh(1, 4, 2, False, True, 0)
    a, b, c, d = y.addVertex(-8, -8, 0), y.addVertex(8, -8, 0), y.addVertex(8, 8, 0), y.addVertex(-8, 8, 0)
    u0, u1, u2, u3 = y.addUv(0.0, 0.0), y.addUv(1.0, 0.0), y.addUv(1.0, 1.0), y.addUv(0.0, 1.0)
    y.addTriangle(a, b, c, u0, u1, u2, "ground"); y.addTriangle(a, c, d, u0, u2, u3, "ground")
    y.endTriMesh()
    for i, (m, orco, uv) in enumerate(box_mats):
        x0, y0 = -3.3 + 2.3 * (i % 3), -0.6 + 2.5 * (i // 3)
        box(2 + i, (x0, y0, 0.003), (x0 + 1.5, y0 + 1.5, 1.1 + 0.2 * i), m, orco, uv, rot=13.0 * i)
    y.endGeometry()
    y.paramsClearAll(); y.paramsSetString("type", "pointlight"); y.paramsSetPoint("from", -2.0, -5.0, 6.0)
    y.paramsSetColor("color", 1.0, 0.95, 0.9); y.paramsSetFloat("power", 60.0); y.createLight("L")
    if integ.get("area_light"):
        y.paramsClearAll(); y.paramsSetString("type", "arealight"); y.paramsSetPoint("corner", 2.0, -2.0, 5.0); y.paramsSetPoint("point1", 3.0, -2.0, 5.0)
        y.paramsSetPoint("point2", 2.0, -1.0, 5.0); y.paramsSetColor("color", 0.8, 0.9, 1.0); y.paramsSetFloat("power", 25.0); y.paramsSetInt("samples", 2); y.createLight("A")
    y.paramsClearAll(); y.paramsSetString("type", "perspective")
    y.paramsSetPoint("from", 0.5, -7.5, 4.5); y.paramsSetPoint("to", 0.0, 1.0, 0.6); y.paramsSetPoint("up", 0.5, -7.5, 5.5)
    y.paramsSetInt("resx", width); y.paramsSetInt("resy", height); y.paramsSetFloat("focal", 1.3)
    _set_params(y, integ.get("camera", {}))   # e.g. {"type": "orthographic", "scale": 9.0}: the other camera types under the same view
    y.createCamera("cam")
    y.paramsClearAll(); y.paramsSetString("type", "constant"); y.paramsSetColor("color", 0.3, 0.4, 0.6); y.paramsSetFloat("power", 0.7); y.createBackground("bg")
    y.paramsClearAll(); y.paramsSetString("type", integrator); y.paramsSetInt("raydepth", integ.get("raydepth", 3)); y.paramsSetInt("shadowDepth", 3)
    if integrator == "pathtracing":
        y.paramsSetInt("path_samples", integ.get("path_samples", 2)); y.paramsSetInt("bounces", integ.get("bounces", 4))
        y.paramsSetInt("russian_roulette_min_bounces", integ.get("rr_min_bounces", 99))
    y.createIntegrator("pt")
    scenes._vol_none(y)
    scenes.render_block(y, width, height, spp)
    return y


# Beer absorption inside glass / rough glass (the volume handler the factories create when the material has a "name" parameter)
EXACT["zoo_dl_beer"] = lambda: material_zoo_scene(96, 54, 2, integrator="directlighting", raydepth=5, beer=True)
EXACT["zoo_pt_beer_norr"] = lambda: material_zoo_scene(96, 54, 4, sun=False, ibl=False, rr_min_bounces=99, beer=True, raydepth=5)
# bgPortalLight: a window mesh that radiates the (HDR / gradient) background
EXACT["zoo_dl_portal_hdr"] = lambda: material_zoo_scene(96, 54, 2, integrator="directlighting", raydepth=3, sun=False, meshlight=False, ibl=False, portal=True,
                                                        hdr_env={"rotation": 15.0})
EXACT["zoo_pt_portal_norr"] = lambda: material_zoo_scene(96, 54, 4, sun=False, meshlight=False, ibl=False, portal=True, rr_min_bounces=99, gradient={
    "horizon_color": (0.9, 0.8, 0.7, 1.0), "zenith_color": (0.2, 0.4, 0.9, 1.0), "power": 1.5})
EXACT["tex_dl_basic"] = lambda: textured_scene(96, 54, 2)
EXACT["tex_dl_modes_b"] = lambda: textured_scene(96, 54, 2, variant="modes_b", tex={"interpolate": "none", "clipping": "extend"})
EXACT["tex_dl_modes_c"] = lambda: textured_scene(96, 54, 2, variant="modes_c", tex={"interpolate": "bicubic", "mirror_x": True, "mirror_y": True, "xrepeat": 2, "yrepeat": 3})
EXACT["tex_dl_adjust"] = lambda: textured_scene(96, 54, 2, tex={"adj_contrast": 1.3, "adj_intensity": 0.9, "adj_saturation": 1.4, "adj_hue": 40.0, "adj_clamp": True,
                                                                "adj_mult_factor_red": 0.9, "adj_mult_factor_blue": 1.2, "rot90": True, "cropmin_x": 0.1, "cropmax_x": 0.8, "cropmax_y": 0.9})
EXACT["tex_dl_checker"] = lambda: textured_scene(96, 54, 2, tex={"clipping": "checker", "even_tiles": True, "odd_tiles": False, "checker_dist": 0.2, "xrepeat": 2})
EXACT["tex_dl_clip"] = lambda: textured_scene(96, 54, 2, tex={"clipping": "clip", "cropmin_y": 0.2})
EXACT["tex_pt_norr"] = lambda: textured_scene(96, 54, 4, integrator="pathtracing")
EXACT["tex_dl_camera_coords"] = lambda: textured_scene(96, 54, 2, variant="camera_coords")
# texco "window" = Camera::screenproject of the other four camera types (camera_orthographic.cc:65, camera_architect.cc:68, camera_angular.cc:123, camera_equirectangular.cc:92)
EXACT["tex_dl_window_ortho"] = lambda: textured_scene(96, 54, 2, variant="camera_coords", camera={"type": "orthographic", "scale": 9.0})
EXACT["tex_dl_window_architect"] = lambda: textured_scene(96, 54, 2, variant="camera_coords", camera={"type": "architect", "focal": 1.1})
EXACT["tex_dl_window_angular"] = lambda: textured_scene(96, 54, 2, variant="camera_coords", camera={"type": "angular", "angle": 60.0, "max_angle": 60.0, "circular": False})
EXACT["tex_dl_window_equirect"] = lambda: textured_scene(96, 48, 2, variant="camera_coords", camera={"type": "equirectangular"})
EXACT["tex_dl_mask"] = lambda: textured_scene(96, 54, 2, variant="mask", raydepth=4)
EXACT["tex_pt_mask_norr"] = lambda: textured_scene(96, 54, 4, variant="mask", integrator="pathtracing", raydepth=4)
EXACT["tex_dl_procedural"] = lambda: textured_scene(96, 54, 2, variant="procedural")
EXACT["tex_pt_procedural_norr"] = lambda: textured_scene(96, 54, 4, variant="procedural", integrator="pathtracing")
EXACT["tex_dl_noise"] = lambda: textured_scene(96, 54, 2, variant="noise")
EXACT["tex_dl_noise2"] = lambda: textured_scene(96, 54, 2, variant="noise2")
EXACT["tex_pt_noise2_norr"] = lambda: textured_scene(96, 54, 4, variant="noise2", integrator="pathtracing")
EXACT["tex_dl_ramp"] = lambda: textured_scene(96, 54, 2, variant="ramp")
EXACT["tex_pt_noise_norr"] = lambda: textured_scene(96, 54, 4, variant="noise", integrator="pathtracing")
EXACT["tex_dl_materials"] = lambda: textured_scene(96, 54, 2, variant="materials", raydepth=4)
EXACT["tex_pt_materials_norr"] = lambda: textured_scene(96, 54, 4, variant="materials", integrator="pathtracing", raydepth=4)


def film_io_scene(aa_passes, film_save_load, width=32, height=18, inc=3):
    """Cornell box for the YAF_FILMv1 tests: `aa_passes` passes (2 samples, then `inc` per pass, no adaptive threshold), computer node 3,
    base sampling offset 17, the given film_save_load mode."""
    y = scenes.cornell_box(width, height, 2, path_samples=2, bounces=5, raydepth=5, rr_min_bounces=99)   # no roulette: no rand()-seeded prng
    if y.render_params is None:
        y.setRenderParams()
    rp = y.render_params
    rp["AA_passes"], rp["AA_inc_samples"], rp["AA_threshold"] = ("int", aa_passes), ("int", inc), ("float", 0.0)
    rp["adv_computer_node"], rp["adv_base_sampling_offset"] = ("int", 3), ("int", 17)
    rp["film_save_load"] = ("string", film_save_load)
    return y


# three AA passes without adaptive threshold on computer node 3 with a base sampling offset
EXACT["cornell_pt_3pass_node3"] = lambda: film_io_scene(3, "none")


# ieslight (light_ies.cc, util_ies.h): synthetic photometric webs written by hand (tests/golden/*.ies) — a full-sphere type C web with nine
# horizontal planes, and an axially symmetric downlight (one horizontal angle, TILT=INCLUDE block, 70 degree cut-off)
IES_WEB = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "golden", "web_typeC.ies")
IES_DOWN = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "golden", "downlight_axial.ies")
EXACT["zoo_dl_ies"] = lambda: material_zoo_scene(96, 54, 2, extra_lights=[
    {"type": "ieslight", "file": IES_WEB, "from": (-1.0, -1.5, 3.5), "to": (-0.6, 0.4, 0.0), "color": (1.0, 0.95, 0.85, 1.0), "power": 30.0},
    {"type": "ieslight", "file": IES_DOWN, "from": (2.2, 0.5, 3.0), "to": (2.0, 0.8, 0.0), "color": (0.6, 0.8, 1.0, 1.0), "power": 25.0}], **_ZOO_DL)
EXACT["zoo_dl_ies_soft"] = lambda: material_zoo_scene(96, 54, 2, extra_lights=[
    {"type": "ieslight", "file": IES_WEB, "from": (0.5, -2.0, 4.0), "to": (0.2, 0.5, 0.0), "soft_shadows": True, "samples": 3, "color": (1.0, 0.9, 0.8, 1.0), "power": 30.0}], **_ZOO_DL)
# path tracing picks ONE light per vertex from a schedule-dependent counter (integrator_montecarlo.cc:62-76): exact parity needs a single light
EXACT["zoo_pt_ies_norr"] = lambda: material_zoo_scene(96, 54, 4, sun=False, ibl=False, meshlight=False, rr_min_bounces=99, extra_lights=[
    {"type": "ieslight", "file": IES_WEB, "from": (-1.0, -2.0, 3.0), "to": (0.0, 0.0, 0.5), "soft_shadows": True, "samples": 2, "color": (0.9, 0.9, 1.0, 1.0), "power": 30.0}])
EXACT["zoo_pt_ies_down_norr"] = lambda: material_zoo_scene(96, 54, 4, sun=False, ibl=False, meshlight=False, rr_min_bounces=99, extra_lights=[
    {"type": "ieslight", "file": IES_DOWN, "from": (0.5, 0.0, 3.2), "to": (0.4, 0.3, 0.0), "color": (1.0, 0.9, 0.8, 1.0), "power": 30.0}])
