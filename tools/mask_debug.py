"""Development tool: variants of the mask parity scene against the live oracle, to isolate which ingredient differs."""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libyafaray_b200 import Context, render_scene, film_to_image
from tests import oracle_util as ou, parity_cases as pc

ctx = Context(0)
orig = pc._textured_mask_materials


def variant(mb3):
    def f(y, sd, node, mapper, layer):
        def mat(typ, **kw):
            y.paramsClearAll(); y.paramsSetString("type", typ); pc._set_params(y, kw)
        real_create = y.createMaterial
        def create(name):
            if name == "mb3":
                return None
            return real_create(name)
        y.createMaterial = create
        orig(y, sd, node, mapper, layer)
        y.createMaterial = real_create
        mat("mask_mat", **mb3["mat"])
        mapper("k3_m", texco="orco", mapping=mb3.get("mapping", "sphere"))
        layer("k3_l", "k3_m", do_color=False, do_scalar=True, color_input=True, def_val=1.0, upper_value=0.0, negative=mb3.get("negative", True))
        y.createMaterial("mb3")
    return f


base = dict(material1="mk_l", material2="mk_a", threshold=0.5, mask="k3_l", receive_shadows=False)
V = {
    "as_is": dict(mat=base),
    "shadows_on": dict(mat=dict(base, receive_shadows=True)),
    "swapped": dict(mat=dict(base, material1="mk_a", material2="mk_l")),
    "both_a": dict(mat=dict(base, material1="mk_a", material2="mk_a")),
    "both_l": dict(mat=dict(base, material1="mk_l", material2="mk_l")),
    "cube": dict(mat=base, mapping="cube"),
    "positive": dict(mat=base, negative=False),
}
for name, v in V.items():
    for rd in (0, 4):
        pc._textured_mask_materials = variant(v)
        y = pc.textured_scene(96, 54, 2, variant="mask", raydepth=rd)
        g = film_to_image(render_scene(y, ctx=ctx))
        rf, _ = ou.run_reference_render(y)
        r = film_to_image(rf)
        d = np.abs(g[..., :3] - r[..., :3]).max(axis=2)
        top = d[:24, :40]
        print(json.dumps({"variant": name, "raydepth": rd, "nbad_box3": int((top > 2e-5).sum()), "max_box3": float(top.max()), "nbad_rest": int((d > 2e-5).sum() - (top > 2e-5).sum())}))
        sys.stdout.flush()
