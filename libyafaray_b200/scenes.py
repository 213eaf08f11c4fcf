"""Synthetic workloads of BASELINE.json's configs, built through the Interface mirror so that the SAME description
goes to the CUDA path (scene_desc) and, as XML, to the reference (oracle).  Shapes follow SURVEY.md §8(d).
"""
import numpy as np

from .interface import Interface


def _quad(y, a, b, c, d, mat):
    """two triangles a-b-c, a-c-d"""
    ia, ib, ic, id_ = [y.addVertex(*p) for p in (a, b, c, d)]
    y.addTriangle(ia, ib, ic, mat)
    y.addTriangle(ia, ic, id_, mat)


def _box(y, lo, hi, mat, rot_deg=0.0):
    """axis-aligned box rotated about its vertical axis; outward-facing triangles (12)"""
    cx, cy = 0.5 * (lo[0] + hi[0]), 0.5 * (lo[1] + hi[1])
    cs, sn = np.cos(np.radians(rot_deg)), np.sin(np.radians(rot_deg))

    def P(x, yy, z):
        dx, dy = x - cx, yy - cy
        return (float(cx + cs * dx - sn * dy), float(cy + sn * dx + cs * dy), float(z))
    v = [P(lo[0], lo[1], lo[2]), P(hi[0], lo[1], lo[2]), P(hi[0], hi[1], lo[2]), P(lo[0], hi[1], lo[2]),
         P(lo[0], lo[1], hi[2]), P(hi[0], lo[1], hi[2]), P(hi[0], hi[1], hi[2]), P(lo[0], hi[1], hi[2])]
    idx = [y.addVertex(*p) for p in v]
    faces = [(0, 2, 1), (0, 3, 2), (4, 5, 6), (4, 6, 7), (0, 1, 5), (0, 5, 4), (1, 2, 6), (1, 6, 5), (2, 3, 7), (2, 7, 6),
             (3, 0, 4), (3, 4, 7)]
    for a, b, c in faces:
        y.addTriangle(idx[a], idx[b], idx[c], mat)


def _shiny(y, name, col, **kw):
    y.paramsClearAll()
    y.paramsSetString("type", "shinydiffusemat")
    y.paramsSetColor("color", *col)
    for k, v in kw.items():
        if isinstance(v, bool): y.paramsSetBool(k, v)
        elif isinstance(v, int): y.paramsSetInt(k, v)
        elif isinstance(v, float): y.paramsSetFloat(k, v)
        elif isinstance(v, str): y.paramsSetString(k, v)
        else: y.paramsSetColor(k, *v)
    return y.createMaterial(name)


def render_block(y, width, height, spp, integrator="pt", camera="cam", background="bg", aa_passes=1, filter_type="box",
                 pixelwidth=1.5, threads=1, inc_samples=None, **extra):
    y.paramsClearAll()
    y.paramsSetString("camera_name", camera)
    y.paramsSetString("integrator_name", integrator)
    y.paramsSetString("volintegrator_name", "volintegr")
    if background:
        y.paramsSetString("background_name", background)
    y.paramsSetInt("width", width); y.paramsSetInt("height", height)
    y.paramsSetInt("AA_passes", aa_passes); y.paramsSetInt("AA_minsamples", spp)
    y.paramsSetInt("AA_inc_samples", inc_samples if inc_samples is not None else spp)
    y.paramsSetFloat("AA_threshold", 0.0)
    y.paramsSetFloat("AA_pixelwidth", pixelwidth)
    y.paramsSetString("filter_type", filter_type)
    y.paramsSetString("color_space", "LinearRGB")
    y.paramsSetString("tiles_order", "linear")
    y.paramsSetInt("tile_size", 32)
    y.paramsSetInt("threads", threads)
    for k, v in extra.items():
        if isinstance(v, bool): y.paramsSetBool(k, v)
        elif isinstance(v, int): y.paramsSetInt(k, v)
        elif isinstance(v, float): y.paramsSetFloat(k, v)
        else: y.paramsSetString(k, v)
    y.setRenderParams()


def _vol_none(y):
    y.paramsClearAll()
    y.paramsSetString("type", "none")
    y.createIntegrator("volintegr")


def cornell_box(width=1920, height=1080, spp=256, path_samples=1, bounces=8, raydepth=8, rr_min_bounces=2,
                light_samples=4, integrator="pathtracing", glass=True, light_power=30.0, filter_type="box",
                aa_passes=1, extra_point_light=False, transp_shad=False, no_recursive=False, lens=None, dispersion=0.0, rough_glass=False, **render_extra):
    """BASELINE config 2: Cornell box, PathIntegrator, area light, ShinyDiffuse + Glass (36 triangles)."""
    y = Interface()
    y.startScene(0)
    white = _shiny(y, "white", (0.75, 0.75, 0.75))
    red = _shiny(y, "red", (0.63, 0.065, 0.05))
    green = _shiny(y, "green", (0.14, 0.45, 0.09))
    y.paramsClearAll()
    if glass:
        y.paramsSetString("type", "rough_glass" if rough_glass else "glass"); y.paramsSetFloat("IOR", 1.5)
        y.paramsSetColor("filter_color", 1.0, 1.0, 1.0); y.paramsSetColor("mirror_color", 1.0, 1.0, 1.0)
        if rough_glass:
            y.paramsSetFloat("alpha", 0.3)
        if dispersion > 0:
            y.paramsSetFloat("dispersion_power", dispersion)
        gl = y.createMaterial("glassmat")
    else:
        gl = _shiny(y, "glassmat", (0.3, 0.3, 0.7))
    y.paramsClearAll()
    y.paramsSetString("type", "light_mat"); y.paramsSetColor("color", 1.0, 1.0, 1.0); y.paramsSetFloat("power", light_power)
    lm = y.createMaterial("lightmat")

    y.startGeometry()
    y.startTriMesh(1, 20, 10, False, False, 0)
    _quad(y, (-1, -1, 0), (1, -1, 0), (1, 1, 0), (-1, 1, 0), white)       # floor (normal +z)
    _quad(y, (-1, -1, 2), (-1, 1, 2), (1, 1, 2), (1, -1, 2), white)       # ceiling (normal -z)
    _quad(y, (-1, 1, 0), (1, 1, 0), (1, 1, 2), (-1, 1, 2), white)         # back wall
    _quad(y, (-1, -1, 0), (-1, 1, 0), (-1, 1, 2), (-1, -1, 2), red)       # left wall
    _quad(y, (1, -1, 0), (1, -1, 2), (1, 1, 2), (1, 1, 0), green)         # right wall
    y.endTriMesh()
    y.startTriMesh(2, 8, 12, False, False, 0)
    _box(y, (-0.65, 0.05, 0.002), (-0.05, 0.65, 1.2), white, rot_deg=18.0)   # tall diffuse box
    y.endTriMesh()
    y.startTriMesh(3, 8, 12, False, False, 0)
    _box(y, (0.1, -0.55, 0.002), (0.7, 0.05, 0.6), gl, rot_deg=-17.0)   # lifted off the floor: coplanar faces are epsilon ties        # short glass box
    y.endTriMesh()
    y.startTriMesh(20, 4, 2, False, False, 0)
    z = 1.98
    _quad(y, (-0.15, -0.15, z), (-0.15, 0.15, z), (0.15, 0.15, z), (0.15, -0.15, z), lm)  # emitter quad, normal -z
    y.endTriMesh()
    y.endGeometry()

    y.paramsClearAll()
    y.paramsSetString("type", "arealight")
    y.paramsSetPoint("corner", -0.15, -0.15, z)
    y.paramsSetPoint("point1", -0.15, 0.15, z)
    y.paramsSetPoint("point2", 0.15, -0.15, z)
    y.paramsSetColor("color", 1.0, 1.0, 1.0); y.paramsSetFloat("power", light_power)
    y.paramsSetInt("samples", light_samples); y.paramsSetInt("object", 20)
    y.createLight("ceiling")
    if extra_point_light:
        y.paramsClearAll()
        y.paramsSetString("type", "pointlight"); y.paramsSetPoint("from", 0.5, -0.6, 1.5)
        y.paramsSetColor("color", 1.0, 0.9, 0.8); y.paramsSetFloat("power", 1.5)
        y.createLight("fill")

    y.paramsClearAll()
    y.paramsSetString("type", "perspective")
    y.paramsSetPoint("from", 0.0, -3.9, 1.0); y.paramsSetPoint("to", 0.0, 0.0, 1.0); y.paramsSetPoint("up", 0.0, -3.9, 2.0)
    y.paramsSetInt("resx", width); y.paramsSetInt("resy", height); y.paramsSetFloat("focal", 1.1)
    for k, v in (lens or {}).items():   # thin-lens parameters: aperture, dof_distance, bokeh_type, bokeh_bias, bokeh_rotation
        if isinstance(v, str): y.paramsSetString(k, v)
        else: y.paramsSetFloat(k, float(v))
    y.createCamera("cam")

    y.paramsClearAll()
    y.paramsSetString("type", "constant"); y.paramsSetColor("color", 0.02, 0.02, 0.03); y.paramsSetFloat("power", 1.0)
    y.createBackground("bg")

    y.paramsClearAll()
    y.paramsSetString("type", integrator)
    y.paramsSetInt("raydepth", raydepth); y.paramsSetInt("shadowDepth", 4); y.paramsSetBool("transpShad", transp_shad)
    if integrator == "pathtracing":
        y.paramsSetInt("path_samples", path_samples); y.paramsSetInt("bounces", bounces)
        y.paramsSetInt("russian_roulette_min_bounces", rr_min_bounces); y.paramsSetBool("no_recursive", no_recursive)
    y.createIntegrator("pt")
    _vol_none(y)
    render_block(y, width, height, spp, filter_type=filter_type, aa_passes=aa_passes, **render_extra)
    return y


def simple_point_scene(width=160, height=90, spp=1, integrator="directlighting", filter_type="gauss", bg_transp=False, bg_transp_refract=False, see_through=False,
                       ground_quad=True, aa_passes=1, **render_extra):
    """Small DirectLight + point-light scene in the spirit of BASELINE config 1 (tests/test01 geometry is 6 cubes on
    a ground quad lit by one point light; its textures are 'next' scope, so flat colours are used)."""
    y = Interface()
    y.startScene(0)
    ground = _shiny(y, "ground", (0.8, 0.8, 0.8))
    cols = [(0.9, 0.2, 0.2), (0.2, 0.9, 0.2), (0.2, 0.2, 0.9), (0.9, 0.9, 0.2), (0.9, 0.2, 0.9), (0.2, 0.9, 0.9)]
    mats = [_shiny(y, "cube%d" % i, c, diffuse_reflect=0.5 + 0.1 * (i % 3)) for i, c in enumerate(cols)]
    if see_through:
        # materials whose Material::getAlpha is not 1 (the "bg_transp_refract" chain): glass, tinted glass, rough glass, a transparent shinydiffusemat,
        # and a blend of the last with an opaque one
        def glass(name, kind="glass", filt=(1.0, 1.0, 1.0), **kw):
            y.paramsClearAll(); y.paramsSetString("type", kind); y.paramsSetFloat("IOR", 1.5)
            y.paramsSetColor("filter_color", *filt); y.paramsSetColor("mirror_color", 1.0, 1.0, 1.0)
            for k, v in kw.items():
                y.paramsSetFloat(k, v)
            return y.createMaterial(name)
        mats[0] = glass("glass0")
        mats[1] = glass("glass1", filt=(0.9, 0.5, 0.3), transmit_filter=0.8)
        mats[2] = glass("rglass2", "rough_glass", alpha=0.2)
        mats[3] = _shiny(y, "transp3", (0.9, 0.9, 0.2), transparency=0.7, transmit_filter=0.5, IOR=1.3, fresnel_effect=True, specular_reflect=0.2)
        y.paramsClearAll(); y.paramsSetString("type", "blend_mat"); y.paramsSetString("material1", "transp3"); y.paramsSetString("material2", "cube5")
        y.paramsSetFloat("blend_value", 0.4)
        mats[4] = y.createMaterial("blend4")
    y.startGeometry()
    for i in range(6):
        y.startTriMesh(1 + i, 8, 12, False, False, 0)
        x0, y0 = -4.4 + 2.6 * (i % 3), 1.4 - 3.4 * (i // 3)
        _box(y, (x0, y0, 1e-5), (x0 + 2.0, y0 + 2.0, 2.0), mats[i], rot_deg=7.0 * i)
        y.endTriMesh()
    if ground_quad:
        y.startTriMesh(7, 4, 2, False, False, 0)
        _quad(y, (-10, -10, 0), (10, -10, 0), (10, 10, 0), (-10, 10, 0), ground)
        y.endTriMesh()
    y.endGeometry()
    y.paramsClearAll()
    y.paramsSetString("type", "pointlight"); y.paramsSetPoint("from", 5.27648, -4.88993, 8.89514)
    y.paramsSetColor("color", 1.0, 1.0, 1.0); y.paramsSetFloat("power", 72.0)
    y.createLight("Point")
    y.paramsClearAll()
    y.paramsSetString("type", "perspective")
    y.paramsSetPoint("from", 8.64791, -7.22615, 8.1295); y.paramsSetPoint("to", 8.03447, -6.65603, 7.58301)
    y.paramsSetPoint("up", 8.25644, -6.8447, 8.9669)
    y.paramsSetInt("resx", width); y.paramsSetInt("resy", height); y.paramsSetFloat("focal", 1.09375)
    y.createCamera("cam")
    y.paramsClearAll()
    y.paramsSetString("type", "constant"); y.paramsSetColor("color", 0.854296, 0.854296, 0.854296); y.paramsSetFloat("power", 0.5)
    y.createBackground("bg")
    y.paramsClearAll()
    y.paramsSetString("type", integrator); y.paramsSetInt("raydepth", 8); y.paramsSetInt("shadowDepth", 2)
    if bg_transp:
        y.paramsSetBool("bg_transp", True)
    if bg_transp_refract:
        y.paramsSetBool("bg_transp_refract", True)
    if integrator == "pathtracing":
        y.paramsSetInt("path_samples", 2); y.paramsSetInt("bounces", 3); y.paramsSetInt("russian_roulette_min_bounces", 8)
    y.createIntegrator("pt")
    _vol_none(y)
    render_block(y, width, height, spp, filter_type=filter_type, aa_passes=aa_passes, **render_extra)
    return y


def procedural_mesh(n_tris=10_000_000, width=1920, height=1080, spp=16, seed=12345, bounces=8, raydepth=8, path_samples=1,
                    rr_min_bounces=2, integrator="pathtracing", smooth=True, ibl_hdr=None, mesh_lights=0, blend=False, crop=None, threads=1):
    """BASELINE config 3: procedural displaced-grid terrain of ~n_tris triangles (2*n^2) with materials cycling over
    glossy / coated_glossy / rough_glass / shinydiffuse in 16x16-quad patches, a sun and an area light, path depth 8.
    Built through the bulk mesh call (numpy arrays); value-noise displacement from a seeded generator.
    `crop` = (xstart, ystart, width, height): render only that window of the width x height camera image (the <render> block's
    xstart/ystart/width/height, environment.cc:456-470)."""
    n = max(2, int(round((n_tris / 2) ** 0.5)))
    rng = np.random.default_rng(seed)
    y = Interface()
    y.startScene(0)

    def mat(name, typ, **kw):
        y.paramsClearAll(); y.paramsSetString("type", typ)
        for k, v in kw.items():
            if isinstance(v, bool): y.paramsSetBool(k, v)
            elif isinstance(v, int): y.paramsSetInt(k, v)
            elif isinstance(v, float): y.paramsSetFloat(k, v)
            elif isinstance(v, str): y.paramsSetString(k, v)
            else: y.paramsSetColor(k, *v)
        return y.createMaterial(name)
    names = [
        mat("gloss_a", "glossy", color=(0.9, 0.7, 0.3), diffuse_color=(0.6, 0.4, 0.2), diffuse_reflect=0.5, glossy_reflect=0.7, exponent=500.0, as_diffuse=False),
        mat("coat_a", "coated_glossy", color=(0.3, 0.5, 0.9), diffuse_color=(0.2, 0.3, 0.6), diffuse_reflect=0.6, glossy_reflect=0.6, exponent=100.0, IOR=1.8, as_diffuse=False),
        mat("rough_a", "rough_glass", IOR=1.5, alpha=0.2, filter_color=(0.9, 1.0, 0.95), transmit_filter=1.0),
        mat("diff_a", "shinydiffusemat", color=(0.7, 0.7, 0.7)),
        mat("gloss_b", "glossy", color=(0.8, 0.8, 0.8), diffuse_color=(0.4, 0.5, 0.3), diffuse_reflect=0.7, glossy_reflect=0.4, exponent=50.0, as_diffuse=True),
        mat("coat_b", "coated_glossy", color=(0.9, 0.3, 0.3), diffuse_color=(0.6, 0.2, 0.2), diffuse_reflect=0.7, glossy_reflect=0.5, exponent=2000.0, IOR=1.8, as_diffuse=False),
    ]
    if blend:
        # BASELINE config 4: blend_mat of glossy / shinydiffuse with blend_value 0.3 / 0.5 / 0.7 in the material cycle
        for k, bv in enumerate((0.3, 0.5, 0.7)):
            y.paramsClearAll(); y.paramsSetString("type", "blend_mat"); y.paramsSetString("material1", "gloss_a" if k != 1 else "gloss_b")
            y.paramsSetString("material2", "diff_a"); y.paramsSetFloat("blend_value", bv)
            names.append(y.createMaterial("blend_%d" % k))
    # value noise: three octaves of bilinearly interpolated random lattices
    xs = np.linspace(0.0, 1.0, n + 1, dtype=np.float32)
    h = np.zeros((n + 1, n + 1), np.float32)
    for cells, amp in ((8, 1.6), (64, 0.35), (512, 0.06)):
        lat = rng.random((cells + 2, cells + 2)).astype(np.float32)
        f = xs * cells
        i0 = np.minimum(f.astype(np.int32), cells); t = (f - i0).astype(np.float32)
        t = t * t * (3 - 2 * t)
        a = lat[i0][:, i0]; b = lat[i0][:, i0 + 1]; c = lat[i0 + 1][:, i0]; d = lat[i0 + 1][:, i0 + 1]
        tx, ty = t[None, :], t[:, None]
        h += amp * ((a * (1 - tx) + b * tx) * (1 - ty) + (c * (1 - tx) + d * tx) * ty)
    X, Y = np.meshgrid(xs * 20 - 10, xs * 20 - 10)
    P = np.stack([X, Y, h], axis=-1).reshape(-1, 3).astype(np.float32)
    j, i = np.meshgrid(np.arange(n, dtype=np.int32), np.arange(n, dtype=np.int32), indexing="ij")
    a = (j * (n + 1) + i).reshape(-1)
    F = np.empty((2 * n * n, 3), np.int32)
    F[0::2] = np.stack([a, a + 1, a + n + 2], axis=1)
    F[1::2] = np.stack([a, a + n + 2, a + n + 1], axis=1)
    patch = ((j // 16) * 7 + (i // 16) * 3).reshape(-1) % len(names)
    per_tri = np.repeat(patch, 2).astype(np.int32)
    normals = None
    if smooth:
        gy, gx = np.gradient(h, 20.0 / n)
        N = np.stack([-gx, -gy, np.ones_like(h)], axis=-1)
        N /= np.linalg.norm(N, axis=-1, keepdims=True)
        normals = N.reshape(-1, 3).astype(np.float32)
    y.startGeometry()
    y.addTriMeshArrays(1, P, F, (names, per_tri), normals=normals)
    lm = mat("lightmat", "light_mat", color=(1.0, 1.0, 1.0), power=40.0)
    y.startTriMesh(2, 4, 2, False, False, 0)
    z = 6.0
    _quad(y, (-2, -2, z), (-2, 2, z), (2, 2, z), (2, -2, z), lm)
    y.endTriMesh()
    for k in range(mesh_lights):
        # config 4: mesh lights of ~10^3 triangles each (a 22 x 23 grid of quads = 1012 triangles), facing down
        gx, gy = 22, 23
        cx, cy, cz = -5.0 + 5.0 * k, 2.0 - 2.0 * k, 5.0 + 0.5 * k
        ux = np.linspace(-0.8, 0.8, gx + 1, dtype=np.float32); uy = np.linspace(-0.8, 0.8, gy + 1, dtype=np.float32)
        LX, LY = np.meshgrid(ux + cx, uy + cy)
        LP = np.stack([LX, LY, np.full_like(LX, cz) + 0.1 * np.sin(3.0 * LX) * np.cos(3.0 * LY)], axis=-1).reshape(-1, 3).astype(np.float32)
        jj, ii = np.meshgrid(np.arange(gy, dtype=np.int32), np.arange(gx, dtype=np.int32), indexing="ij")
        aa = (jj * (gx + 1) + ii).reshape(-1)
        LF = np.empty((2 * gx * gy, 3), np.int32)
        LF[0::2] = np.stack([aa, aa + gx + 2, aa + 1], axis=1)      # wound so that the geometric normal points down (-z)
        LF[1::2] = np.stack([aa, aa + gx + 1, aa + gx + 2], axis=1)
        y.addTriMeshArrays(10 + k, LP, LF, "lightmat")
    y.endGeometry()
    for k in range(mesh_lights):
        y.paramsClearAll(); y.paramsSetString("type", "meshlight"); y.paramsSetInt("object", 10 + k); y.paramsSetColor("color", 1.0, 0.9 - 0.1 * k, 0.7 + 0.1 * k)
        y.paramsSetFloat("power", 12.0); y.paramsSetInt("samples", 1); y.createLight("mesh%d" % k)
    y.paramsClearAll(); y.paramsSetString("type", "arealight"); y.paramsSetPoint("corner", -2, -2, z); y.paramsSetPoint("point1", -2, 2, z)
    y.paramsSetPoint("point2", 2, -2, z); y.paramsSetColor("color", 1.0, 1.0, 1.0); y.paramsSetFloat("power", 40.0)
    y.paramsSetInt("samples", 1); y.paramsSetInt("object", 2); y.createLight("area")
    y.paramsClearAll(); y.paramsSetString("type", "sunlight"); y.paramsSetPoint("direction", 0.5, -0.4, 0.75); y.paramsSetFloat("angle", 1.0)
    y.paramsSetColor("color", 1.0, 0.95, 0.85); y.paramsSetFloat("power", 1.5); y.paramsSetInt("samples", 1); y.createLight("sun")
    y.paramsClearAll(); y.paramsSetString("type", "perspective")
    y.paramsSetPoint("from", 0.0, -13.0, 7.0); y.paramsSetPoint("to", 0.0, 0.0, 0.8); y.paramsSetPoint("up", 0.0, -13.0, 8.0)
    y.paramsSetInt("resx", width); y.paramsSetInt("resy", height); y.paramsSetFloat("focal", 1.4); y.createCamera("cam")
    if ibl_hdr:
        # config 4: equirectangular HDR "textureback" with ibl true, ibl_samples 16
        y.paramsClearAll(); y.paramsSetString("type", "image"); y.paramsSetString("filename", ibl_hdr); y.createTexture("envtex")
        y.paramsClearAll(); y.paramsSetString("type", "textureback"); y.paramsSetString("texture", "envtex"); y.paramsSetString("mapping", "sphere")
        y.paramsSetFloat("power", 1.0); y.paramsSetBool("ibl", True); y.paramsSetInt("ibl_samples", 16)
    else:
        y.paramsClearAll(); y.paramsSetString("type", "constant"); y.paramsSetColor("color", 0.4, 0.5, 0.7); y.paramsSetFloat("power", 1.0)
    y.createBackground("bg")
    y.paramsClearAll(); y.paramsSetString("type", integrator); y.paramsSetInt("raydepth", raydepth); y.paramsSetInt("shadowDepth", 2)
    if integrator == "pathtracing":
        y.paramsSetInt("path_samples", path_samples); y.paramsSetInt("bounces", bounces)
        y.paramsSetInt("russian_roulette_min_bounces", rr_min_bounces)
    y.createIntegrator("pt")
    _vol_none(y)
    if crop is None:
        render_block(y, width, height, spp, threads=threads)
    else:
        render_block(y, crop[2], crop[3], spp, threads=threads, xstart=int(crop[0]), ystart=int(crop[1]))
    return y
