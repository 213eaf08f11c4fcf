"""Host-side mirror of the reference's scene-building API.

`Interface` follows `yafaray4::Interface` (/root/reference/include/interface/interface.h:48-139,
/root/reference/src/interface/interface.cc:121-457): the same imperative calls (startScene, startTriMesh, addVertex,
addTriangle, paramsSet*, create{Material,Light,Camera,Background,Integrator}, render) with the same argument meaning,
and the same parameter keys and in-code defaults the reference's factories use (cited per factory below).

It produces two things from one description:
  * `to_xml(path)`   — the reference's XML scene format (as parsed by /root/reference/src/common/import_xml.cc), so the
                       oracle renders exactly the same scene through the unchanged XML loader;
  * `scene_desc()`   — the flattened POD descriptor of include/yafgpu.h, handed to the CUDA library through the C ABI.
This module never computes a pixel: all rendering happens in libyafgpu.so.
"""
import ctypes as C
import numpy as np

from . import capi
from .capi import f32


class _Mesh:
    def __init__(self, oid, has_orco, has_uv, typ, pass_index):
        self.id, self.has_orco, self.has_uv, self.type, self.pass_index = oid, has_orco, has_uv, typ, pass_index
        self.points, self.normals, self.uvs = [], [], []   # points: interleaved with orco when has_orco
        self.faces, self.face_uv, self.face_mat = [], [], []
        self.np_points = self.np_faces = self.np_normals = None  # bulk path
        self.np_mat = None
        self.smooth_angle = None
        self.normals_exported = False

    def n_tris(self):
        return len(self.np_faces) if self.np_faces is not None else len(self.faces)


class Interface:
    """Python mirror of yafaray4::Interface (see module docstring)."""

    def __init__(self):
        self.params, self.eparams, self._cparams = {}, [], None
        self._cparams = self.params
        self.materials, self.lights, self.cameras, self.backgrounds, self.integrators, self.textures = {}, {}, {}, {}, {}, {}
        self.meshes = {}
        self._cur = None
        self._next_free_id = 2 ** 31 - 1  # scene.cc:42 state_.next_free_id_ = INT_MAX, counting down
        self.render_params = None
        self.scene_type = 0

    # ---- ParamMap building (interface.cc:239-321) -----------------------------------------------------------------
    def paramsSetPoint(self, name, x, y, z): self._cparams[name] = ("point", (float(x), float(y), float(z)))
    def paramsSetString(self, name, s): self._cparams[name] = ("string", str(s))
    def paramsSetBool(self, name, b): self._cparams[name] = ("bool", bool(b))
    def paramsSetInt(self, name, i): self._cparams[name] = ("int", int(i))
    def paramsSetFloat(self, name, f): self._cparams[name] = ("float", float(f))
    def paramsSetColor(self, name, r, g, b, a=1.0): self._cparams[name] = ("color", (float(r), float(g), float(b), float(a)))

    def paramsClearAll(self):
        self.params, self.eparams = {}, []
        self._cparams = self.params

    def paramsStartList(self):
        self.eparams.append({})
        self._cparams = self.eparams[-1]

    def paramsPushList(self):
        self.eparams.append({})
        self._cparams = self.eparams[-1]

    def paramsEndList(self): self._cparams = self.params

    # ---- scene (interface.cc:121-237) -----------------------------------------------------------------------------
    def startScene(self, type=0):
        self.scene_type = type
        return True

    def startGeometry(self): return True
    def endGeometry(self): return True

    def getNextFreeId(self):
        i = self._next_free_id
        self._next_free_id -= 1
        return i

    def startTriMesh(self, id, vertices, triangles, has_orco, has_uv=False, type=0, obj_pass_index=0):
        if id in self.meshes:
            return False
        self._cur = _Mesh(id, bool(has_orco), bool(has_uv), type, obj_pass_index)
        self.meshes[id] = self._cur
        return True

    def endTriMesh(self):
        self._cur = None
        return True

    def addVertex(self, x, y, z, ox=None, oy=None, oz=None):
        m = self._cur
        m.points.append((x, y, z))
        if m.has_orco:
            m.points.append((ox, oy, oz))
            return (len(m.points) - 1) // 2
        return len(m.points) - 1

    def addNormal(self, nx, ny, nz):
        # Scene::addNormal: attaches to the last vertex, marks normals_exported_ (scene.cc:236-262)
        m = self._cur
        m.normals.append((nx, ny, nz))
        m.normals_exported = True

    def addUv(self, u, v):
        self._cur.uvs.append((u, v))
        return len(self._cur.uvs) - 1

    def addTriangle(self, a, b, c, *rest):
        m = self._cur
        if len(rest) == 1:
            mat = rest[0]
            uv = None
        else:
            uv, mat = tuple(rest[:3]), rest[3]
        m.faces.append((a, b, c))
        m.face_uv.append(uv)
        m.face_mat.append(mat)
        return True

    def addTriMeshArrays(self, id, points, faces, mat, normals=None, obj_pass_index=0):
        """Bulk equivalent of startTriMesh/addVertex*/addTriangle*/endTriMesh for large procedural meshes
        (points float32 [nv,3], faces int32 [nt,3], one material name or an int32 [nt] array of indices into
        `mat` when `mat` is a list of names)."""
        m = _Mesh(id, False, False, 0, obj_pass_index)
        m.np_points = np.ascontiguousarray(points, dtype=np.float32)
        m.np_faces = np.ascontiguousarray(faces, dtype=np.int32)
        m.np_normals = None if normals is None else np.ascontiguousarray(normals, dtype=np.float32)
        m.normals_exported = normals is not None
        m.np_mat = mat
        self.meshes[id] = m
        return True

    def addInstance(self, base_object_id, obj_to_world):
        # TriangleObjectInstance::getPrimitives is declared non-const (object_geom_mesh.h:124) and so does not override the const virtual that
        # Scene::update calls (scene.cc:816): the reference counts an instance's triangles but never hands them to the kd-tree builder, which then
        # reads uninitialised pointers (it crashes on every scene with an <instance>).  There is no reference behaviour to reproduce.
        raise NotImplementedError("addInstance: instances do not render in the reference (Scene::update never collects their triangles)")

    def smoothMesh(self, id, angle):
        raise NotImplementedError("smoothMesh: pass explicit vertex normals (addNormal); the reference-side adaptor "
                                  "exports the normals Scene::smoothMesh produced (scene.cc:420-539)")

    # ---- factories (interface.cc:323-372): store the ParamMap under the name ---------------------------------------
    def _take(self):
        p, e = dict(self.params), [dict(x) for x in self.eparams]
        return p, e

    def createMaterial(self, name):
        self.materials[name] = self._take()
        return name

    def createLight(self, name):
        self.lights[name] = self._take()[0]
        return name

    def createCamera(self, name):
        self.cameras[name] = self._take()[0]
        return name

    def createTexture(self, name):
        self.textures[name] = self._take()[0]
        return name

    def createBackground(self, name):
        self.backgrounds[name] = self._take()[0]
        return name

    def createIntegrator(self, name):
        self.integrators[name] = self._take()[0]
        return name

    def setRenderParams(self):
        """What Interface::render does with the current ParamMap (interface.cc:395-404): they become the <render> block."""
        self.render_params = dict(self.params)

    def render(self, output=None):
        from .render import render_scene
        self.setRenderParams()
        return render_scene(self, output)

    # ---- XML export (format parsed by /root/reference/src/common/import_xml.cc:227-776) ---------------------------
    @staticmethod
    def _xml_param(name, tv, indent="\t"):
        t, v = tv
        if t == "point": return '%s<%s x="%r" y="%r" z="%r"/>\n' % (indent, name, v[0], v[1], v[2])
        if t == "color": return '%s<%s r="%r" g="%r" b="%r" a="%r"/>\n' % (indent, name, v[0], v[1], v[2], v[3])
        if t == "string": return '%s<%s sval="%s"/>\n' % (indent, name, v)
        if t == "bool": return '%s<%s bval="%s"/>\n' % (indent, name, "true" if v else "false")
        if t == "int": return '%s<%s ival="%d"/>\n' % (indent, name, v)
        if t == "float": return '%s<%s fval="%r"/>\n' % (indent, name, v)
        raise ValueError(t)

    def to_xml(self, path):
        """Streams the scene to `path`.  Bulk meshes (addTriMeshArrays) are written in chunks, so a 10^7-triangle mesh costs no
        more host memory than the arrays it is made of; floats are printed with 9 significant digits, which round-trips float32
        through the reference's atof() -> float (import_xml.cc:92-121)."""
        if self.render_params is None:
            self.setRenderParams()
        with open(path, "w", buffering=1 << 22) as fh:
            w = fh.write
            w('<?xml version="1.0"?>\n<scene type="%s">\n\n' % ("triangle" if self.scene_type == 0 else "universal"))
            for name, p in self.textures.items():
                w('<texture name="%s">\n' % name)
                for k, tv in p.items(): w(self._xml_param(k, tv))
                w("</texture>\n\n")
            for name, (p, e) in self.materials.items():
                w('<material name="%s">\n' % name)
                for k, tv in p.items(): w(self._xml_param(k, tv))
                for le in e:
                    w("\t<list_element>\n")
                    for k, tv in le.items(): w(self._xml_param(k, tv, "\t\t"))
                    w("\t</list_element>\n")
                w("</material>\n\n")
            for tag, table in (("light", self.lights), ("camera", self.cameras), ("background", self.backgrounds),
                               ("integrator", self.integrators)):
                for name, p in table.items():
                    w('<%s name="%s">\n' % (tag, name))
                    for k, tv in p.items(): w(self._xml_param(k, tv))
                    w("</%s>\n\n" % tag)
            for oid in sorted(self.meshes):
                m = self.meshes[oid]
                if m.np_faces is not None:
                    self._xml_bulk_mesh(w, oid, m)
                    continue
                nv = len(m.points) // 2 if m.has_orco else len(m.points)
                w('<mesh id="%d" vertices="%d" faces="%d" has_orco="%s" has_uv="%s" type="%d" obj_pass_index="%d">\n'
                  % (oid, nv, len(m.faces), "true" if m.has_orco else "false", "true" if m.has_uv else "false", m.type, m.pass_index))
                for i in range(nv):
                    if m.has_orco:
                        p, o = m.points[2 * i], m.points[2 * i + 1]
                        w('\t<p x="%r" y="%r" z="%r" ox="%r" oy="%r" oz="%r"/>\n' % (p[0], p[1], p[2], o[0], o[1], o[2]))
                    else:
                        p = m.points[i]
                        w('\t<p x="%r" y="%r" z="%r"/>\n' % (p[0], p[1], p[2]))
                    if m.normals_exported:
                        n = m.normals[i]
                        w('\t<n x="%r" y="%r" z="%r"/>\n' % (n[0], n[1], n[2]))
                for u, v in m.uvs:
                    w('\t<uv u="%r" v="%r"/>\n' % (u, v))
                cur = None
                for f, uv, mat in zip(m.faces, m.face_uv, m.face_mat):
                    if mat != cur:
                        w('\t<set_material sval="%s"/>\n' % mat)
                        cur = mat
                    if uv is None: w('\t<f a="%d" b="%d" c="%d"/>\n' % f)
                    else: w('\t<f a="%d" b="%d" c="%d" uv_a="%d" uv_b="%d" uv_c="%d"/>\n' % (f + uv))
                w("</mesh>\n\n")
            w("<render>\n")
            for k, tv in self.render_params.items(): w(self._xml_param(k, tv))
            w("</render>\n</scene>\n")
        return path

    @staticmethod
    def _xml_bulk_mesh(w, oid, m, chunk=1 << 16):
        """A mesh given as arrays, in the element order the reference's mesh state expects (import_xml.cc:401-416 opens it,
        :550-630 reads `p` [+ `n` for the vertex just added, Scene::addNormal scene.cc:236-262], `set_material`, `f`)."""
        P, F, N = m.np_points, m.np_faces, m.np_normals
        nv, nt = len(P), len(F)
        w('<mesh id="%d" vertices="%d" faces="%d" has_orco="false" has_uv="false" type="%d" obj_pass_index="%d">\n' % (oid, nv, nt, m.type, m.pass_index))
        for s in range(0, nv, chunk):
            if N is None:
                w("".join(['\t<p x="%.9g" y="%.9g" z="%.9g"/>\n' % tuple(r) for r in P[s:s + chunk].tolist()]))
            else:
                rows = np.concatenate([P[s:s + chunk], N[s:s + chunk]], axis=1).tolist()
                w("".join(['\t<p x="%.9g" y="%.9g" z="%.9g"/>\n\t<n x="%.9g" y="%.9g" z="%.9g"/>\n' % tuple(r) for r in rows]))
        if isinstance(m.np_mat, str):
            names, per_tri = [m.np_mat], np.zeros(nt, np.int32)
        else:
            names, per_tri = m.np_mat
            per_tri = np.asarray(per_tri)
        # runs of equal material: one <set_material> per run
        starts = np.concatenate([[0], np.flatnonzero(per_tri[1:] != per_tri[:-1]) + 1, [nt]]) if nt else np.array([0, 0])
        buf, pending = [], 0
        for k in range(len(starts) - 1):
            a, b = int(starts[k]), int(starts[k + 1])
            if b <= a: continue
            buf.append('\t<set_material sval="%s"/>\n' % names[int(per_tri[a])])
            buf.extend(['\t<f a="%d" b="%d" c="%d"/>\n' % tuple(r) for r in F[a:b].tolist()])
            pending += b - a
            if pending >= chunk:
                w("".join(buf)); buf, pending = [], 0
        w("".join(buf))
        w("</mesh>\n\n")

    # ---- flattening to the C ABI descriptor ----------------------------------------------------------------------
    @staticmethod
    def _g(p, key, default):
        """ParamMap::getParam semantics: missing key keeps the in-code default (include/common/param.h:91-106)."""
        if key not in p:
            return default
        return p[key][1]

    def _material_record(self, name, index_of):
        p, _e = self.materials[name]
        g = self._g
        t = g(p, "type", "")
        m = capi.Material()
        vis = {"normal": capi.VIS_NORMAL, "no_shadows": capi.VIS_NO_SHADOWS, "shadow_only": capi.VIS_SHADOW_ONLY,
               "invisible": capi.VIS_INVISIBLE}.get(g(p, "visibility", "normal"), capi.VIS_NORMAL)
        m.visibility = vis
        m.receive_shadows = int(g(p, "receive_shadows", True))
        m.flat_material = int(g(p, "flat_material", False))
        m.additional_depth = int(g(p, "additionaldepth", 0))
        m.transparent_bias_factor = g(p, "transparentbias_factor", 0.0)
        m.transparent_bias_multiply_raydepth = int(g(p, "transparentbias_multiply_raydepth", False))
        m.sampling_factor = g(p, "samplingfactor", 1.0)
        m.mat_pass_index = int(g(p, "mat_pass_index", 0))
        self._node_records(name, m)
        col = lambda key, d: capi.F3(*[np.float32(x) for x in g(p, key, d)[:3]])
        if t == "shinydiffusemat":  # material_shiny_diffuse.cc:599-688
            m.type = capi.MAT_SHINYDIFFUSE
            s = m.u.shinydiffuse
            s.color, s.mirror_color = col("color", (1, 1, 1, 1)), col("mirror_color", (1, 1, 1, 1))
            s.diffuse_reflect, s.specular_reflect = g(p, "diffuse_reflect", 1.0), g(p, "specular_reflect", 0.0)
            s.transparency, s.translucency, s.emit = g(p, "transparency", 0.0), g(p, "translucency", 0.0), g(p, "emit", 0.0)
            s.ior, s.fresnel_effect = g(p, "IOR", 1.33), int(g(p, "fresnel_effect", False))
            s.transmit_filter = g(p, "transmit_filter", 1.0)
            s.oren_nayar = int(g(p, "diffuse_brdf", "") == "oren_nayar")
            s.sigma = g(p, "sigma", 0.1)
        elif t in ("glass", "rough_glass"):  # material_glass.cc:311-379, material_rough_glass.cc:322-401
            m.type = capi.MAT_GLASS if t == "glass" else capi.MAT_ROUGH_GLASS
            s = m.u.glass
            s.ior = g(p, "IOR", 1.4)
            s.filter_color, s.mirror_color = col("filter_color", (1, 1, 1, 1)), col("mirror_color", (1, 1, 1, 1))
            s.transmit_filter, s.dispersion_power = g(p, "transmit_filter", 0.0), g(p, "dispersion_power", 0.0)
            s.fake_shadows = int(g(p, "fake_shadows", False))
            s.alpha = g(p, "alpha", 0.5)
            ab = g(p, "absorption", (1, 1, 1, 1))
            # the Beer volume handler only exists when the material's own ParamMap has a "name" (material_glass.cc:390-398)
            s.has_absorption = int("absorption" in p and min(ab[:3]) < 1.0 and "name" in p)
            s.absorption = capi.F3(*[np.float32(x) for x in ab[:3]])
            s.absorption_dist = g(p, "absorption_dist", 1.0)
        elif t in ("glossy", "coated_glossy"):  # material_glossy.cc:407-480, material_coated_glossy.cc:465-545
            m.type = capi.MAT_GLOSSY if t == "glossy" else capi.MAT_COATED_GLOSSY
            s = m.u.glossy
            s.color, s.diffuse_color = col("color", (1, 1, 1, 1)), col("diffuse_color", (1, 1, 1, 1))
            s.mirror_color = col("mirror_color", (1, 1, 1, 1))
            s.diffuse_reflect, s.glossy_reflect = g(p, "diffuse_reflect", 0.0), g(p, "glossy_reflect", 1.0)
            s.specular_reflect = g(p, "specular_reflect", 1.0)
            s.exponent, s.exp_u, s.exp_v = g(p, "exponent", 50.0), g(p, "exp_u", 50.0), g(p, "exp_v", 50.0)
            s.ior = g(p, "IOR", 1.4)
            s.as_diffuse, s.anisotropic = int(g(p, "as_diffuse", True)), int(g(p, "anisotropic", False))
            s.oren_nayar = int(g(p, "diffuse_brdf", "") == "Oren-Nayar")
            s.sigma = g(p, "sigma", 0.1)
        elif t == "mask_mat":  # material_mask.cc:150-198
            m.type = capi.MAT_MASK
            # the mask factory reads nothing but threshold, the two children, receive_shadows and visibility: the rest stays at Material's defaults
            m.flat_material, m.additional_depth, m.transparent_bias_factor, m.transparent_bias_multiply_raydepth = 0, 0, 0.0, 0
            m.sampling_factor, m.mat_pass_index = 1.0, 0
            m.u.blend.material1 = index_of(g(p, "material1", ""))
            m.u.blend.material2 = index_of(g(p, "material2", ""))
            m.u.blend.blend_value = g(p, "threshold", 0.5)
        elif t == "blend_mat":  # material_blend.cc:464-518
            m.type = capi.MAT_BLEND
            m.u.blend.material1 = index_of(g(p, "material1", ""))
            m.u.blend.material2 = index_of(g(p, "material2", ""))
            m.u.blend.blend_value = g(p, "blend_value", 0.5)
        elif t == "light_mat":  # material_simple.cc:63-73
            m.type = capi.MAT_LIGHT
            m.u.simple.color = col("color", (1, 1, 1, 1))
            m.u.simple.strength = g(p, "power", 1.0)
            m.u.simple.double_sided = int(g(p, "double_sided", False))
        elif t == "null":
            m.type = capi.MAT_NULL
        else:
            raise NotImplementedError("material type '%s'" % t)
        return m

    # ---- shader nodes and textures (material_node.cc:140-237, shader_node_basic.cc:344-420, shader_node_layer.cc:160-244) ----------
    def _node_records(self, name, m):
        """Appends this material's <list_element element="shader_node"> entries to self._nodes (declaration order) and binds the
        "<slot>_shader" parameters, exactly what NodeMaterial::loadNodes / parseNodes read.  A texture_mapper whose texture does not
        exist makes loadNodes fail in the reference, which then leaves the material without any node: mirrored."""
        p, elems = self.materials[name]
        g = self._g
        for s in range(capi.N_SLOTS):
            m.shader[s] = -1
        m.node_first, m.node_count = len(self._nodes), 0
        elems = [e for e in elems if g(e, "element", "shader_node") == "shader_node"]
        if not elems:
            return
        names = [g(e, "name", None) for e in elems]
        if None in names or len(set(names)) != len(names):
            return  # loadNodes error -> no nodes
        idx = {n: i for i, n in enumerate(names)}
        recs = []
        for e in elems:
            t = g(e, "type", None)
            r = capi.ShaderNode()
            r.input1 = r.input2 = r.factor = r.texture = -1

            def ref(key):
                if key not in e:
                    return -1
                n = g(e, key, "")
                if n not in idx:
                    raise KeyError("shader node '%s': input '%s' does not exist" % (g(e, "name", ""), n))
                return idx[n]
            col4 = lambda key, d: (f32 * 4)(*[np.float32(x) for x in g(e, key, d)])
            if t == "texture_mapper":
                r.type = capi.NODE_TEXTURE_MAPPER
                tn = g(e, "texture", None)
                if tn is None or tn not in self.textures:
                    return  # "texture does not exist" -> factory returns nullptr -> loadNodes fails -> the material has no nodes
                r.texture = self._texture_index(tn)
                r.texco = capi.TEXCO.get(g(e, "texco", "global"), capi.TEXCO["global"])
                r.mapping = capi.TEXPROJ.get(g(e, "mapping", "plain"), 0)
                r.proj_x, r.proj_y, r.proj_z = int(g(e, "proj_x", 1)), int(g(e, "proj_y", 2)), int(g(e, "proj_z", 3))
                r.scale = capi.F3(*[np.float32(x) for x in g(e, "scale", (1, 1, 1))])
                r.offset = capi.F3(*[np.float32(x) for x in g(e, "offset", (0, 0, 0))])
                r.bump_strength, r.do_scalar = g(e, "bump_strength", 1.0), int(g(e, "do_scalar", True))
                if "transform" in e:
                    raise NotImplementedError("texture_mapper 'transform' matrices are not passed by this host mirror")
                r.transform = (f32 * 16)(1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1)
            elif t == "value":
                r.type = capi.NODE_VALUE
                c = g(e, "color", (1, 1, 1, 1))
                r.color = (f32 * 4)(np.float32(c[0]), np.float32(c[1]), np.float32(c[2]), np.float32(g(e, "alpha", 1.0)))
                r.scalar = g(e, "scalar", 1.0)
            elif t == "mix":
                r.type, r.mode = capi.NODE_MIX, int(g(e, "mode", 0))
                r.input1, r.input2, r.factor = ref("input1"), ref("input2"), ref("factor")
                if (r.input1 < 0 and "color1" not in e) or (r.input2 < 0 and "color2" not in e) or (r.factor < 0 and "value" not in e):
                    return  # configInputs fails -> no nodes
                r.color, r.color2 = col4("color1", (0, 0, 0, 1)), col4("color2", (0, 0, 0, 1))
                r.scalar = g(e, "value", 0.0)
            elif t == "layer":
                r.type, r.mode = capi.NODE_LAYER, int(g(e, "mode", 0))
                r.input1, r.input2 = ref("input"), ref("upper_layer")
                if r.input1 < 0:
                    return
                dc = g(e, "def_col", (1, 1, 1, 1))
                r.color = (f32 * 4)(np.float32(dc[0]), np.float32(dc[1]), np.float32(dc[2]), 1.0)
                r.color2 = col4("upper_color", (0, 0, 0, 1))
                r.scalar, r.scalar2 = g(e, "def_val", 1.0), g(e, "upper_value", 0.0)
                r.colfac, r.valfac = g(e, "colfac", 1.0), g(e, "valfac", 1.0)
                r.do_color, r.layer_do_scalar = int(g(e, "do_color", True)), int(g(e, "do_scalar", False))
                r.color_input, r.use_alpha = int(g(e, "color_input", True)), int(g(e, "use_alpha", False))
                r.no_rgb, r.stencil, r.negative = int(g(e, "noRGB", False)), int(g(e, "stencil", False)), int(g(e, "negative", False))
            else:
                raise NotImplementedError("shader node type '%s'" % t)
            recs.append(r)
        mtype = g(p, "type", "")
        for key, slot in capi.SLOTS.items():
            if key == "mask" and mtype not in ("blend_mat", "mask_mat"):
                continue
            if key in p:
                n = g(p, key, "")
                if n in idx:   # parseNodes only warns about a missing node (material_node.cc:216-235)
                    m.shader[slot] = idx[n]
        self._nodes += recs
        m.node_count = len(recs)

    def _color_ramp(self, t, p):
        """textureReadColorRamp__ (texture_basic.cc:54-90) + ColorRamp's constructor and addItem (color_ramp.cc:31-56): items sorted by position."""
        g = self._g
        n = int(g(p, "ramp_num_items", 0))
        if not g(p, "use_color_ramp", False) or n <= 0:
            return
        items = []
        for i in range(n):
            col = g(p, "ramp_item_%d_color" % i, (0.0, 0.0, 0.0, 1.0))
            items.append((np.float32(g(p, "ramp_item_%d_position" % i, 0.0)), [np.float32(col[0]), np.float32(col[1]), np.float32(col[2]), np.float32(g(p, "ramp_item_%d_alpha" % i, 1.0))]))
        items.sort(key=lambda it: it[0])
        arr = np.array([c + [pos] for pos, c in items], np.float32)
        self._tex_keep.append(arr)
        t.ramp_items, t.ramp = n, arr.ctypes.data_as(C.POINTER(C.c_float))
        # the constructor accepts exactly the upper- and the lower-case spelling of each keyword
        t.ramp_mode = {"HSV": 1, "hsv": 1, "HSL": 2, "hsl": 2}.get(g(p, "ramp_color_mode", ""), 0)
        t.ramp_interpolation = 0 if g(p, "ramp_interpolation", "") in ("CONSTANT", "constant") else 1
        t.ramp_hue_interpolation = {"FAR": 1, "far": 1, "CW": 2, "cw": 2, "CCW": 3, "ccw": 3}.get(g(p, "ramp_hue_interpolation", ""), 0)

    def _texture_index(self, name):
        if name in self._tex_index:
            return self._tex_index[name]
        p, g = self.textures[name], self._g
        ttype = g(p, "type", "")
        if ttype in ("blend", "rgb_cube"):   # texture_basic.cc:759-776, 363-386
            t = capi.Texture()
            t.adj_intensity = t.adj_contrast = t.adj_saturation = t.adj_mult_red = t.adj_mult_green = t.adj_mult_blue = 1.0
            if ttype == "blend":
                self._color_ramp(t, p)
                t.type, t.blend_type = capi.TEX_BLEND, capi.TEXBLEND.get(g(p, "stype", ""), 0)   # unknown strings: linear
                t.use_flip_axis = int(g(p, "use_flip_axis", False))
            else:
                t.type = capi.TEX_RGB_CUBE
                t.adj_intensity, t.adj_contrast = g(p, "adj_intensity", 1.0), g(p, "adj_contrast", 1.0)
                t.adj_saturation, t.adj_hue = g(p, "adj_saturation", 1.0), float(np.float32(g(p, "adj_hue", 0.0)) / np.float32(60.0))
                t.adj_mult_red, t.adj_mult_green, t.adj_mult_blue = g(p, "adj_mult_factor_red", 1.0), g(p, "adj_mult_factor_green", 1.0), g(p, "adj_mult_factor_blue", 1.0)
                t.adj_clamp = int(g(p, "adj_clamp", False))
            self._tex_index[name] = len(self._texs)
            self._texs.append(t)
            return self._tex_index[name]
        if ttype in ("voronoi", "musgrave", "distorted_noise"):   # texture_basic.cc:456-503, 553-596, 635-673 (factories)
            t = capi.Texture()
            self._color_ramp(t, p)
            t.type = {"voronoi": capi.TEX_VORONOI, "musgrave": capi.TEX_MUSGRAVE, "distorted_noise": capi.TEX_DISTORTED_NOISE}[ttype]
            noise = lambda nt: capi.NOISE.get(nt, 3 if "voronoi" in nt else 0)
            c1, c2 = g(p, "color1", (0.0, 0.0, 0.0)), g(p, "color2", (1.0, 1.0, 1.0))
            for k in range(3):
                t.color1[k], t.color2[k] = c1[k], c2[k]
            t.noise_size = g(p, "size", 1.0)
            if ttype == "voronoi":
                for k, dflt in enumerate((1.0, 0.0, 0.0, 0.0)):
                    t.vor_weight[k] = g(p, "weight%d" % (k + 1), dflt)
                t.mk_exponent = g(p, "mk_exponent", 2.5)
                aw = [np.float32(abs(np.float32(t.vor_weight[k]))) for k in range(4)]
                s = np.float32(np.float32(np.float32(aw[0] + aw[1]) + aw[2]) + aw[3])     # iscale_ = aw_1_ + aw_2_ + aw_3_ + aw_4_ (texture_basic.cc:417-418)
                t.intensity_scale = float(np.float32(g(p, "intensity", 1.0)) / s) if s != 0 else 0.0
                t.distance_metric = capi.DIST_METRIC.get(g(p, "distance_metric", ""), 0)
                t.color_type = {"col1": 1, "col2": 2, "col3": 3}.get(g(p, "Rgbype", ""), 0)   # the key the factory reads (texture_basic.cc:473)
            elif ttype == "musgrave":
                t.noise_type, t.musgrave_type = noise(g(p, "noise_type", "")), capi.MUSGRAVE.get(g(p, "musgrave_type", ""), 0)
                t.mus_h, t.mus_lacunarity, t.mus_octaves = g(p, "H", 1.0), g(p, "lacunarity", 2.0), g(p, "octaves", 2.0)
                t.mus_offset, t.mus_gain, t.intensity_scale = g(p, "offset", 1.0), g(p, "gain", 1.0), g(p, "intensity", 1.0)
            else:
                t.noise_type, t.noise_type2, t.distort = noise(g(p, "noise_type1", "")), noise(g(p, "noise_type2", "")), g(p, "distort", 1.0)
            t.adj_intensity, t.adj_contrast = g(p, "adj_intensity", 1.0), g(p, "adj_contrast", 1.0)
            t.adj_saturation, t.adj_hue = g(p, "adj_saturation", 1.0), float(np.float32(g(p, "adj_hue", 0.0)) / np.float32(60.0))
            t.adj_mult_red, t.adj_mult_green, t.adj_mult_blue = g(p, "adj_mult_factor_red", 1.0), g(p, "adj_mult_factor_green", 1.0), g(p, "adj_mult_factor_blue", 1.0)
            t.adj_clamp = int(g(p, "adj_clamp", False))
            self._uses_noise = True
            self._tex_index[name] = len(self._texs)
            self._texs.append(t)
            return self._tex_index[name]
        if ttype in ("clouds", "marble", "wood"):   # texture_basic.cc:96-171, 177-250, 257-343 (factories), newNoise__ :24-52
            t = capi.Texture()
            self._color_ramp(t, p)
            t.type = {"clouds": capi.TEX_CLOUDS, "marble": capi.TEX_MARBLE, "wood": capi.TEX_WOOD}[ttype]
            nt = g(p, "noise_type", "")
            t.noise_type = capi.NOISE.get(nt, 3 if "voronoi" in nt else 0)
            t.noise_depth, t.noise_size, t.noise_hard = int(g(p, "depth", 2)), g(p, "size", 1.0), int(g(p, "hard", False))
            c1, c2 = g(p, "color1", (0.0, 0.0, 0.0)), g(p, "color2", (1.0, 1.0, 1.0))
            for k in range(3):
                t.color1[k], t.color2[k] = c1[k], c2[k]
            t.clouds_bias = {"positive": 1, "negative": 2}.get(g(p, "bias", ""), 0) if ttype == "clouds" else 0
            t.turbulence = g(p, "turbulence", 1.0)
            shp = np.float32(g(p, "sharpness", 1.0))
            t.sharpness = float(np.float32(1.0 / float(shp))) if (ttype == "marble" and shp > 1) else 1.0
            t.wave_shape = {"saw": 1, "tri": 2}.get(g(p, "shape", ""), 0)
            t.wood_rings = int(g(p, "wood_type", "") == "rings")
            t.adj_intensity, t.adj_contrast = g(p, "adj_intensity", 1.0), g(p, "adj_contrast", 1.0)
            t.adj_saturation, t.adj_hue = g(p, "adj_saturation", 1.0), float(np.float32(g(p, "adj_hue", 0.0)) / np.float32(60.0))
            t.adj_mult_red, t.adj_mult_green, t.adj_mult_blue = g(p, "adj_mult_factor_red", 1.0), g(p, "adj_mult_factor_green", 1.0), g(p, "adj_mult_factor_blue", 1.0)
            t.adj_clamp = int(g(p, "adj_clamp", False))
            self._uses_noise = True
            self._tex_index[name] = len(self._texs)
            self._texs.append(t)
            return self._tex_index[name]
        if ttype != "image":
            raise NotImplementedError("texture type '%s' (image, blend and rgb_cube textures are implemented on the device)" % ttype)
        fn = g(p, "filename", "")
        if isinstance(fn, str) and fn.lower().endswith(".hdr"):
            rgb = read_radiance_hdr(fn)   # HDR: always linear, never "optimized" (texture_image.cc:600-606)
            cs = capi.CS["LinearRGB"]
        else:
            raise NotImplementedError("image texture '%s': only Radiance .hdr is decoded by this host mirror (the C++ adaptor "
                                      "takes any format the reference's image handlers load)" % fn)
        rgba = np.ones(rgb.shape[:2] + (4,), np.float32)
        rgba[..., :3] = rgb
        rgba = np.ascontiguousarray(rgba)
        t = capi.Texture()
        t.type, t.width, t.height, t.rgba = capi.TEX_IMAGE, rgba.shape[1], rgba.shape[0], rgba.ctypes.data_as(capi.pf)
        interp = g(p, "interpolate", "bilinear")
        if interp in ("mipmap_trilinear", "mipmap_ewa"):
            raise NotImplementedError("texture interpolate '%s' (ray differentials) is not implemented on the device" % interp)
        t.interpolate = capi.TEXINTERP.get(interp, 1)
        t.clip_mode = capi.TEXCLIP.get(g(p, "clipping", "repeat"), capi.TEXCLIP["repeat"])  # string2Cliptype__: unknown -> repeat
        t.xrepeat, t.yrepeat = int(g(p, "xrepeat", 1)), int(g(p, "yrepeat", 1))
        t.mirror_x, t.mirror_y, t.rot90 = int(g(p, "mirror_x", False)), int(g(p, "mirror_y", False)), int(g(p, "rot90", False))
        t.cropmin_x, t.cropmin_y = g(p, "cropmin_x", 0.0), g(p, "cropmin_y", 0.0)
        t.cropmax_x, t.cropmax_y = g(p, "cropmax_x", 1.0), g(p, "cropmax_y", 1.0)
        t.checker_even, t.checker_odd = int(g(p, "even_tiles", False)), int(g(p, "odd_tiles", True))
        t.checker_dist = g(p, "checker_dist", 0.0)
        t.normalmap, t.color_space, t.gamma = int(g(p, "normalmap", False)), cs, g(p, "gamma", 1.0)
        t.adj_intensity, t.adj_contrast = g(p, "adj_intensity", 1.0), g(p, "adj_contrast", 1.0)
        t.adj_saturation, t.adj_hue = g(p, "adj_saturation", 1.0), float(np.float32(g(p, "adj_hue", 0.0)) / np.float32(60.0))
        t.adj_mult_red, t.adj_mult_green, t.adj_mult_blue = g(p, "adj_mult_factor_red", 1.0), g(p, "adj_mult_factor_green", 1.0), g(p, "adj_mult_factor_blue", 1.0)
        t.adj_clamp = int(g(p, "adj_clamp", False))
        self._tex_index[name] = len(self._texs)
        self._texs.append(t)
        self._tex_keep.append(rgba)
        return self._tex_index[name]

    def _light_record(self, p, obj_index_of_id):
        g = self._g
        t = g(p, "type", "")
        l = capi.Light()
        l.light_enabled, l.cast_shadows = int(g(p, "light_enabled", True)), int(g(p, "cast_shadows", True))
        l.photon_only = int(g(p, "photon_only", False))
        l.shoot_caustic, l.shoot_diffuse = int(g(p, "with_caustic", True)), int(g(p, "with_diffuse", True))
        l.color = capi.F3(*[np.float32(x) for x in g(p, "color", (1, 1, 1, 1))[:3]])
        l.power = g(p, "power", 1.0)
        l.samples = int(g(p, "samples", 4))
        l.object = -1
        pt = lambda key, d: capi.F3(*[np.float32(x) for x in g(p, key, d)])
        if t == "pointlight":  # light_point.cc:101-133
            l.type, l.from_ = capi.LIGHT_POINT, pt("from", (0, 0, 0))
        elif t == "sunlight":  # light_sun.cc:95-125
            l.type, l.from_, l.angle = capi.LIGHT_SUN, pt("direction", (0, 0, 1)), g(p, "angle", 0.27)
        elif t == "arealight":  # light_area.cc:174-213
            l.type = capi.LIGHT_AREA
            l.from_, l.point1, l.point2 = pt("corner", (0, 0, 0)), pt("point1", (0, 0, 0)), pt("point2", (0, 0, 0))
            l.object = obj_index_of_id(int(g(p, "object", 0)))
        elif t == "meshlight":  # light_meshlight.cc:228-263
            l.type = capi.LIGHT_MESH
            l.object = obj_index_of_id(int(g(p, "object", 0)))
            l.double_sided = int(g(p, "double_sided", False))
        elif t == "bgPortalLight":  # light_background_portal.cc:236-266
            l.type = capi.LIGHT_BGPORTAL
            l.object = obj_index_of_id(int(g(p, "object", 0)))
        elif t == "directional":  # light_directional.cc:112-158
            l.type, l.point1, l.infinite = capi.LIGHT_DIRECTIONAL, pt("direction", (0, 0, 1)), int(g(p, "infinite", True))
            if not l.infinite:
                l.from_ = pt("from", g(p, "position", (0, 0, 0)))
                l.radius = g(p, "radius", 1.0)
        elif t == "spotlight":  # light_spot.cc:251-295
            l.type, l.from_, l.point1 = capi.LIGHT_SPOT, pt("from", (0, 0, 0)), pt("to", (0, 0, -1))
            l.cone_angle, l.blend = g(p, "cone_angle", 45.0), g(p, "blend", 0.15)
            l.soft_shadows, l.shadow_fuzzy, l.samples = int(g(p, "soft_shadows", False)), g(p, "shadowFuzzyness", 1.0), int(g(p, "samples", 8))
        elif t == "ieslight":  # light_ies.cc:184-233; the library parses the file contents as IesData::parseIesFile does
            l.type, l.from_, l.point1 = capi.LIGHT_IES, pt("from", (0, 0, 0)), pt("to", (0, 0, -1))
            l.soft_shadows, l.samples = int(g(p, "soft_shadows", False)), int(g(p, "samples", 16))
            with open(g(p, "file", ""), "rb") as fh:   # a missing file: the reference's factory returns no light and the scene load fails
                data = fh.read()
            self._ies_keep.append(data)
            l.ies_data, l.ies_size = data, len(data)
        elif t == "spherelight":  # light_sphere.cc:163-198
            l.type, l.from_, l.radius = capi.LIGHT_SPHERE, pt("from", (0, 0, 0)), g(p, "radius", 1.0)
            if int(g(p, "object", 0)) != 0:
                raise NotImplementedError("spherelight linked to a sphere primitive ('object')")
        else:
            raise NotImplementedError("light type '%s'" % t)
        return l

    def _background_texels(self, name):
        """Texels of the image texture `name` as ImageHandler::getPixel(x, y) delivers them, float32 [h, w, 3].
        Only Radiance .hdr files (always linear float storage, texture_image.cc:600-606) with the factory's default mapping
        (clipping "repeat", no repeat/mirror/crop/rot90/adjustments) are taken; everything else is refused."""
        if name not in self.textures:
            raise KeyError("unknown texture '%s'" % name)
        p, g = self.textures[name], self._g
        if g(p, "type", "") != "image":
            raise NotImplementedError("texture type '%s' as a background" % g(p, "type", ""))
        fn = g(p, "filename", "")
        if not fn.lower().endswith(".hdr"):
            raise NotImplementedError("background image '%s': only Radiance .hdr is read by this host mirror" % fn)
        if g(p, "interpolate", "bilinear") not in ("bilinear", "none"):
            raise NotImplementedError("texture interpolate '%s'" % g(p, "interpolate", ""))
        defaults = {"xrepeat": 1, "yrepeat": 1, "cropmin_x": 0.0, "cropmin_y": 0.0, "cropmax_x": 1.0, "cropmax_y": 1.0, "rot90": False,
                    "mirror_x": False, "mirror_y": False, "clipping": "repeat", "adj_mult_factor_red": 1.0, "adj_mult_factor_green": 1.0,
                    "adj_mult_factor_blue": 1.0, "adj_intensity": 1.0, "adj_contrast": 1.0, "adj_saturation": 1.0, "adj_hue": 0.0, "adj_clamp": False}
        for k, d in defaults.items():
            if g(p, k, d) != d:
                raise NotImplementedError("texture parameter %s=%r is not implemented on the device" % (k, g(p, k, d)))
        return read_radiance_hdr(fn)

    def scene_desc(self):
        """Flatten to capi.SceneDesc.  Returns (desc, keepalive) — keepalive owns the numpy/ctypes buffers."""
        if self.render_params is None:
            self.setRenderParams()
        g, rp = self._g, self.render_params
        keep = []
        mat_names = list(self.materials.keys())
        mat_index = {n: i for i, n in enumerate(mat_names)}

        def index_of(n):
            if n not in mat_index:
                raise KeyError("unknown material '%s'" % n)
            return mat_index[n]
        mats = (capi.Material * max(1, len(mat_names)))()
        self._nodes, self._texs, self._tex_index, self._tex_keep, self._uses_noise = [], [], {}, [], False
        for i, n in enumerate(mat_names):
            mats[i] = self._material_record(n, index_of)

        # geometry in the reference's global triangle order: ascending object id (scene.cc:796-817)
        pts, nrm, uvs, vidx, nidx, uvidx, tmat, tobj, objs = [], [], [], [], [], [], [], [], []
        p_off = n_off = uv_off = t_off = 0
        any_n = any_uv = False
        obj_index = {}
        for oid in sorted(self.meshes):
            m = self.meshes[oid]
            if (m.type & 0xFF) != 0:
                raise NotImplementedError("only TRIM meshes (type 0) are supported")
            if m.type & 0x0100 or m.type & 0x0200:  # INVISIBLEM / BASEMESH are skipped by Scene::update (scene.cc:799-800)
                continue
            o = capi.Object()
            o.has_orco, o.has_uv, o.smooth, o.light = int(m.has_orco), int(m.has_uv), int(m.normals_exported), -1
            o.first_tri, o.obj_id, o.pass_index = t_off, oid & 0x7fffffff, m.pass_index
            obj_index[oid] = len(objs)
            if m.np_faces is not None:
                P, F = m.np_points, m.np_faces
                nt = len(F)
                pts.append(P)
                vidx.append(F + p_off)
                if m.np_normals is not None:
                    nrm.append(m.np_normals); nidx.append(F + n_off); n_off += len(m.np_normals); any_n = True
                else:
                    nidx.append(np.full((nt, 3), -1, np.int32))
                uvidx.append(np.full((nt, 3), -1, np.int32))
                if isinstance(m.np_mat, str):
                    tmat.append(np.full(nt, index_of(m.np_mat), np.int32))
                else:
                    names, per_tri = m.np_mat
                    lut = np.array([index_of(n) for n in names], np.int32)
                    tmat.append(lut[np.asarray(per_tri)])
                p_off += len(P)
            else:
                P = np.array(m.points, np.float32).reshape(-1, 3)
                F = np.array(m.faces, np.int32).reshape(-1, 3)
                nt = len(F)
                stride = 2 if m.has_orco else 1  # Scene::addTriangle doubles the indices with orco (scene.cc:615)
                pts.append(P)
                vidx.append(F * stride + p_off)
                if m.normals_exported:
                    # normals_ is resized to points_.size() and indexed by vertex id (scene.cc:594-604,619-633)
                    N = np.zeros((len(P), 3), np.float32)
                    nn = np.array(m.normals, np.float32).reshape(-1, 3)
                    N[:len(nn)] = nn
                    nrm.append(N); nidx.append(F + n_off); n_off += len(N); any_n = True
                else:
                    nidx.append(np.full((nt, 3), -1, np.int32))
                if m.has_uv:
                    UV = np.array(m.uvs, np.float32).reshape(-1, 2)
                    uvs.append(UV)
                    uvidx.append(np.array([u for u in m.face_uv], np.int32).reshape(-1, 3) + uv_off)
                    uv_off += len(UV); any_uv = True
                else:
                    uvidx.append(np.full((nt, 3), -1, np.int32))
                tmat.append(np.array([index_of(n) for n in m.face_mat], np.int32))
                p_off += len(P)
            tobj.append(np.full(nt, len(objs), np.int32))
            o.n_tris = nt
            t_off += nt
            objs.append(o)

        def cat(lst, shape, dt):
            a = np.concatenate(lst).astype(dt, copy=False) if lst else np.zeros(shape, dt)
            a = np.ascontiguousarray(a)
            keep.append(a)
            return a
        A_pts, A_nrm, A_uv = cat(pts, (0, 3), np.float32), cat(nrm, (0, 3), np.float32), cat(uvs, (0, 2), np.float32)
        A_v, A_n, A_u = cat(vidx, (0, 3), np.int32), cat(nidx, (0, 3), np.int32), cat(uvidx, (0, 3), np.int32)
        A_m, A_o = cat(tmat, (0,), np.int32), cat(tobj, (0,), np.int32)
        objarr = (capi.Object * max(1, len(objs)))(*objs)
        pfl = lambda a: a.ctypes.data_as(capi.pf)
        pin = lambda a: a.ctypes.data_as(capi.pi)
        d = capi.SceneDesc()
        d.abi_version = capi.ABI_VERSION
        ge = d.geometry
        ge.points, ge.n_points = pfl(A_pts), len(A_pts)
        ge.normals, ge.n_normals = pfl(A_nrm), len(A_nrm)
        ge.uv, ge.n_uv = pfl(A_uv), len(A_uv)
        ge.tri_vidx, ge.tri_mat, ge.tri_obj, ge.n_tris = pin(A_v), pin(A_m), pin(A_o), len(A_m)
        ge.tri_nidx = pin(A_n) if any_n else None
        ge.tri_uvidx = pin(A_u) if any_uv else None
        ge.objects, ge.n_objects = objarr, len(objs)
        d.materials, d.n_materials = mats, len(mat_names)
        narr = (capi.ShaderNode * max(1, len(self._nodes)))(*self._nodes)
        tarr = (capi.Texture * max(1, len(self._texs)))(*self._texs)
        d.nodes, d.n_nodes, d.textures, d.n_textures = narr, len(self._nodes), tarr, len(self._texs)
        keep += [narr, tarr, self._tex_keep]

        def obj_index_of_id(oid):
            return obj_index.get(oid, -1) if oid else -1
        self._ies_keep = []
        lts = [self._light_record(p, obj_index_of_id) for p in self.lights.values()]
        keep.append(self._ies_keep)
        larr = (capi.Light * max(1, len(lts)))(*lts)
        d.lights, d.n_lights = larr, len(lts)
        keep += [mats, objarr, larr]

        # cameras (camera_perspective.cc:206-252, camera_architect.cc:100-133, camera_orthographic.cc:82-106,
        # camera_angular.cc:100-133, camera_equirectangular.cc:82-108)
        cp = self.cameras[g(rp, "camera_name", next(iter(self.cameras)))]
        ctype = g(cp, "type", "")
        cam_types = {"perspective": capi.CAM_PERSPECTIVE, "architect": capi.CAM_ARCHITECT, "orthographic": capi.CAM_ORTHOGRAPHIC,
                     "angular": capi.CAM_ANGULAR, "equirectangular": capi.CAM_EQUIRECTANGULAR}
        if ctype not in cam_types:
            raise NotImplementedError("camera type '%s'" % ctype)
        c = d.camera
        c.type, c.resx, c.resy = cam_types[ctype], int(g(cp, "resx", 320)), int(g(cp, "resy", 200))
        c.from_ = capi.F3(*[np.float32(x) for x in g(cp, "from", (0, 1, 0))])
        c.to = capi.F3(*[np.float32(x) for x in g(cp, "to", (0, 0, 0))])
        c.up = capi.F3(*[np.float32(x) for x in g(cp, "up", (0, 1, 1))])
        c.focal, c.aperture, c.dof_distance = g(cp, "focal", 1.0), g(cp, "aperture", 0.0), g(cp, "dof_distance", 0.0)
        c.aspect_ratio, c.near_clip = g(cp, "aspect_ratio", 1.0), g(cp, "nearClip", 0.0)
        c.far_clip = g(cp, "farClip", -1.0e38 if ctype in ("angular", "equirectangular") else -1.0)
        c.bokeh_type, c.bokeh_bias = capi.BK.get(g(cp, "bokeh_type", "disk1"), 0), capi.BB.get(g(cp, "bokeh_bias", "uniform"), 0)
        c.bokeh_rotation = g(cp, "bokeh_rotation", 0.0)
        c.scale = g(cp, "scale", 1.0)
        c.angle = float(g(cp, "angle", 90.0))
        c.max_angle = float(g(cp, "max_angle", c.angle))
        c.circular, c.mirrored = int(g(cp, "circular", True)), int(g(cp, "mirrored", False))
        c.projection = capi.ANGULAR_PROJECTIONS.get(g(cp, "projection", ""), 0)

        # background (background_constant.cc:52-86)
        b = d.background
        b.type = capi.BG_NONE
        bname = g(rp, "background_name", None)
        if bname is not None and bname in self.backgrounds:
            bp = self.backgrounds[bname]
            btype = g(bp, "type", "")
            if btype == "constant":
                b.type = capi.BG_CONSTANT
                b.color = capi.F3(*[np.float32(x) for x in g(bp, "color", (0, 0, 0, 1))[:3]])
            elif btype == "textureback":  # background_texture.cc:90-165
                b.type = capi.BG_TEXTURE
                tex = self._background_texels(g(bp, "texture", ""))
                keep.append(tex)
                b.tex_rgb, b.tex_w, b.tex_h = tex.ctypes.data_as(capi.pf), tex.shape[1], tex.shape[0]
                b.tex_interp = int(g(self.textures[g(bp, "texture", "")], "interpolate", "bilinear") != "none")
                b.mapping_angular = int(g(bp, "mapping", "") in ("probe", "angular"))
                b.rotation = g(bp, "rotation", 0.0)
                if g(bp, "smartibl_blur", 0.0) > 0:
                    raise NotImplementedError("textureback: smartibl_blur (mipmapped IBL) is not implemented on the device")
            elif btype == "gradientback":  # background_gradient.cc:66-117
                b.type = capi.BG_GRADIENT
                col = lambda key, dflt: capi.F3(*[np.float32(x) for x in g(bp, key, dflt)[:3]])
                b.horizon_color, b.zenith_color = col("horizon_color", (1, 1, 1, 1)), col("zenith_color", (0.4, 0.5, 1.0, 1))
                b.horizon_ground_color = col("horizon_ground_color", tuple(b.horizon_color))
                b.zenith_ground_color = col("zenith_ground_color", tuple(b.zenith_color))
            else:
                raise NotImplementedError("background type '%s'" % btype)
            b.power, b.ibl, b.with_caustic = g(bp, "power", 1.0), int(g(bp, "ibl", False)), int(g(bp, "with_caustic", True))
            if b.ibl:
                # ConstantBackground::factory creates and adds a "bglight" AFTER all parsed lights (background_constant.cc:69-83)
                bl = capi.Light()
                bl.type, bl.light_enabled, bl.cast_shadows = capi.LIGHT_BACKGROUND, 1, int(g(bp, "cast_shadows", True))
                bl.samples, bl.object = int(g(bp, "ibl_samples", 16)), -1
                bl.shoot_caustic, bl.shoot_diffuse = int(g(bp, "with_caustic", True)), int(g(bp, "with_diffuse", True))
                bl.abs_intersect, bl.clamp_intersect = 0, (g(bp, "ibl_clamp_sampling", 0.0) if btype == "textureback" else 0.0)
                lts.append(bl)
                larr = (capi.Light * len(lts))(*lts)
                d.lights, d.n_lights = larr, len(lts)
                keep.append(larr)

        # film + AA (environment.cc:456-584, 679-813)
        f = d.film
        f.width, f.height = int(g(rp, "width", 320)), int(g(rp, "height", 240))
        f.xstart, f.ystart = int(g(rp, "xstart", 0)), int(g(rp, "ystart", 0))
        f.aa_pixelwidth = g(rp, "AA_pixelwidth", 1.5)
        f.filter_type = capi.FILTERS.get(g(rp, "filter_type", "box"), 0)
        f.tile_size, f.premult = int(g(rp, "tile_size", 32)), int(g(rp, "premult", False))
        f.aa_clamp_samples, f.aa_clamp_indirect = g(rp, "AA_clamp_samples", 0.0), g(rp, "AA_clamp_indirect", 0.0)
        f.aa_passes = int(g(rp, "AA_passes", 1))
        f.aa_light_sample_multiplier = f.aa_indirect_sample_multiplier = 1.0
        f.base_sampling_offset = int(g(rp, "adv_base_sampling_offset", 0)) + int(g(rp, "adv_computer_node", 0)) * 100000

        # integrator (integrator_path_tracer.cc:349-423, integrator_direct_light.cc:196-250)
        ip = self.integrators[g(rp, "integrator_name", next(iter(self.integrators)))]
        it = d.integrator
        tname = g(ip, "type", "")
        if tname == "pathtracing":
            it.type = capi.INT_PATHTRACING
            it.raydepth, it.shadow_depth = int(g(ip, "raydepth", 5)), int(g(ip, "shadowDepth", 5))
            it.path_samples, it.bounces = int(g(ip, "path_samples", 32)), int(g(ip, "bounces", 3))
            it.rr_min_bounces, it.no_recursive = int(g(ip, "russian_roulette_min_bounces", 0)), int(g(ip, "no_recursive", False))
            cm = g(ip, "caustic_type", "path")
            if cm in ("photon", "both"):
                raise NotImplementedError("caustic photon maps are out of scope (SURVEY.md §8f #4)")
            it.trace_caustics = int(cm != "none")  # ctor default caustic_type_ = Path (integrator_path_tracer.cc:35)
        elif tname == "directlighting":
            it.type = capi.INT_DIRECTLIGHT
            it.raydepth, it.shadow_depth = int(g(ip, "raydepth", 5)), int(g(ip, "shadowDepth", 5))
            if g(ip, "caustics", False):
                raise NotImplementedError("caustic photon maps are out of scope (SURVEY.md §8f #4)")
        else:
            raise NotImplementedError("integrator type '%s'" % tname)
        it.transp_shad = int(g(ip, "transpShad", False))
        it.bg_transp, it.bg_transp_refract = int(g(ip, "bg_transp", False)), int(g(ip, "bg_transp_refract", False))
        it.do_ao, it.ao_samples, it.ao_distance = int(g(ip, "do_AO", False)), int(g(ip, "AO_samples", 32)), g(ip, "AO_distance", 1.0)
        it.ao_color = capi.F3(*[np.float32(x) for x in g(ip, "AO_color", (1, 1, 1, 1))[:3]])
        it.shadow_bias_auto, it.shadow_bias = int(g(rp, "adv_auto_shadow_bias_enabled", True)), g(rp, "adv_shadow_bias_value", 0.0005)
        it.ray_min_dist_auto, it.ray_min_dist = int(g(rp, "adv_auto_min_raydist_enabled", True)), g(rp, "adv_min_raydist_value", 0.00005)
        if any(m.type in (capi.MAT_GLASS, capi.MAT_ROUGH_GLASS) and m.u.glass.dispersion_power > 0 for m in mats):
            # CIE 1931 colour matching functions for wl2RgbFromCie__ (spectrum.cc:276-289): data dumped from the reference by tools/make_cie_table.py
            import os
            cie = np.ascontiguousarray(np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "cie_xyz_360_830.npy")), np.float32).reshape(-1, 3)
            keep.append(cie)
            d.cie_xyz, d.n_cie_rows = cie.ctypes.data_as(C.POINTER(C.c_float)), len(cie)
        if self._uses_noise:
            # lattice tables of the noise bases (noise_generator.cc:25-320), dumped from the reference's source by tools/make_noise_tables.py
            import os
            nt = np.fromfile(os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "noise_tables.bin"), np.uint8)
            keep.append(nt)
            d.noise_tables, d.noise_tables_bytes = nt.ctypes.data, nt.nbytes
        return d, keep


def read_radiance_hdr(path):
    """Radiance RGBE reader matching the reference's HdrHandler (src/imagehandler/imagehandler_hdr.cc, imagehandler_util_hdr.h:88-100):
    texel = mantissa * 2^(e - 136) (0 when e == 0), standard "-Y h +X w" orientation, flat or new-RLE scanlines.  float32 [h, w, 3]."""
    with open(path, "rb") as fh:
        data = fh.read()
    pos, w, h = 0, None, None
    while True:
        end = data.index(b"\n", pos)
        line = data[pos:end].decode("latin-1").strip()
        pos = end + 1
        if line.startswith("-Y") or line.startswith("+Y"):
            t = line.split()
            if t[0] != "-Y" or t[2] != "+X":
                raise NotImplementedError("hdr orientation '%s'" % line)
            h, w = int(t[1]), int(t[3])
            break
    buf = np.frombuffer(data, np.uint8, offset=pos)
    out = np.zeros((h, w, 4), np.uint8)
    o = 0
    for y in range(h):
        if 8 <= w <= 0x7fff and buf[o] == 2 and buf[o + 1] == 2 and (int(buf[o + 2]) << 8 | int(buf[o + 3])) == w:
            o += 4
            for ch in range(4):
                x = 0
                while x < w:
                    c = int(buf[o]); o += 1
                    if c > 128:
                        out[y, x:x + c - 128, ch] = buf[o]; o += 1; x += c - 128
                    else:
                        out[y, x:x + c, ch] = buf[o:o + c]; o += c; x += c
        else:
            # flat / old-RLE scanline.  HdrHandler::readOrle copies only the FIRST pixel of such a scanline into the image
            # (its store loop steps by max_[1] instead of step_[1], imagehandler_hdr.cc:283-288); runs are not expanded here
            # because nothing but that pixel is visible to the renderer.  Mirror it: the reference's texels are the definition.
            out[y, 0] = buf[o:o + 4]; o += 4 * w
    e = out[..., 3].astype(np.int32)
    f = np.where(e > 0, np.ldexp(np.float32(1.0), e - 136), np.float32(0.0)).astype(np.float32)
    return np.ascontiguousarray(out[..., :3].astype(np.float32) * f[..., None])


def write_radiance_hdr(path, img):
    """RGBE writer for test fixtures (new-RLE scanline framing); img float32 [h, w, 3]."""
    img = np.asarray(img, np.float32)
    h, w, _ = img.shape
    m = img.max(axis=2)
    mant, ex = np.frexp(m)
    scale = np.where(m > 1e-32, mant * 255.9999 / np.where(m > 1e-32, m, 1), 0).astype(np.float32)
    rgbe = np.zeros((h, w, 4), np.uint8)
    rgbe[..., :3] = (img * scale[..., None]).astype(np.uint8)
    rgbe[..., 3] = np.where(m > 1e-32, ex + 128, 0).astype(np.uint8)
    with open(path, "wb") as fh:
        fh.write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n" % (h, w))
        if not (8 <= w <= 0x7fff):
            fh.write(rgbe.tobytes())
            return
        for y in range(h):   # new-RLE scanline made of literal chunks only (valid, just not compressed)
            fh.write(bytes([2, 2, w >> 8, w & 255]))
            for ch in range(4):
                row = rgbe[y, :, ch].tobytes()
                for x in range(0, w, 128):
                    chunk = row[x:x + 128]
                    fh.write(bytes([len(chunk)])); fh.write(chunk)
