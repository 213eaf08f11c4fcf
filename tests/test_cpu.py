"""CPU-side tests (`-m "not gpu"`): the C ABI library loads and exports every declared symbol, the host mirror of the
reference interface, the oracle and its golden fixtures, and the world_size-2 film reduction (gloo)."""
import ctypes
import os
import re
import subprocess
import sys
import tempfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import oracle_util as ou, parity_cases as pc, qmc_ref  # noqa: E402


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "yafgpu.h")).read()
    names = set(re.findall(r"^(?:int|void|const char \*|void \*|yafgpu_ctx \*)\s*\*?(yafgpu_\w+)\(", hdr, re.M))
    assert len(names) >= 32
    lib = ctypes.CDLL(os.path.join(ROOT, "libyafaray_b200", "libyafgpu.so"))
    for n in names:
        assert hasattr(lib, n), n
    from libyafaray_b200 import capi
    assert set(capi.SYMBOLS) == names
    assert lib.yafgpu_abi_version() == capi.ABI_VERSION


def test_struct_sizes_match_header():
    """ctypes mirrors vs the C compiler's view of include/yafgpu.h."""
    from libyafaray_b200 import capi
    src = '#include "yafgpu.h"\n#include <stdio.h>\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(yafgpu_material), sizeof(yafgpu_light), sizeof(yafgpu_scene_desc), sizeof(yafgpu_pass_desc), sizeof(yafgpu_stats), sizeof(yafgpu_camera), sizeof(yafgpu_object), sizeof(yafgpu_texture), sizeof(yafgpu_shader_node));return 0;}\n'
    with tempfile.TemporaryDirectory() as td:
        open(os.path.join(td, "a.c"), "w").write(src)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), os.path.join(td, "a.c"), "-o", os.path.join(td, "a")])
        out = subprocess.check_output([os.path.join(td, "a")]).decode().split()
    want = [ctypes.sizeof(t) for t in (capi.Material, capi.Light, capi.SceneDesc, capi.PassDesc, capi.Stats, capi.Camera, capi.Object, capi.Texture, capi.ShaderNode)]
    assert [int(x) for x in out] == want


def test_no_gpu_fails_loudly():
    """Without a CUDA device the product must refuse, not fall back."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from libyafaray_b200 import Context, capi
    with pytest.raises(capi.YafGpuError):
        Context(0)


def test_interface_xml_and_descriptor():
    from libyafaray_b200 import scenes
    y = scenes.cornell_box(64, 36, 2)
    d, keep = y.scene_desc()
    assert d.geometry.n_tris == 36 and d.n_lights == 1 and d.n_materials == 5
    objs = [d.geometry.objects[i] for i in range(d.geometry.n_objects)]
    assert [o.first_tri for o in objs] == [0, 10, 22, 34]           # ascending object id = global triangle order
    assert d.lights[0].object == 3                                   # arealight tagged to mesh id 20
    with tempfile.TemporaryDirectory() as td:
        xml = y.to_xml(os.path.join(td, "s.xml"))
        txt = open(xml).read()
    assert txt.count("<f ") == 36 and 'sval="pathtracing"' in txt


def test_descriptor_of_ies_lights_and_procedural_textures():
    """Host mirror only (no device call): an ieslight carries the file CONTENTS across the ABI, blend / rgb_cube textures carry their
    factory parameters, mask_mat keeps only what its factory reads, and what the device refuses is refused here too."""
    from libyafaray_b200 import capi
    y = pc.EXACT["zoo_dl_ies_soft"]()
    d, keep = y.scene_desc()
    ies = [d.lights[i] for i in range(d.n_lights) if d.lights[i].type == capi.LIGHT_IES]
    assert len(ies) == 1 and ies[0].soft_shadows == 1 and ies[0].samples == 3
    data = open(pc.IES_WEB, "rb").read()
    assert ies[0].ies_size == len(data) and C_string(ies[0].ies_data, ies[0].ies_size) == data
    y = pc.EXACT["tex_dl_procedural"]()
    d, keep = y.scene_desc()
    tex = [d.textures[i] for i in range(d.n_textures)]
    kinds = sorted((t.type, t.blend_type, t.use_flip_axis) for t in tex)
    assert (capi.TEX_BLEND, capi.TEXBLEND["radial"], 0) in kinds and (capi.TEX_BLEND, capi.TEXBLEND["diag"], 1) in kinds
    cube = [t for t in tex if t.type == capi.TEX_RGB_CUBE and t.adj_clamp]
    assert len(cube) == 1 and abs(cube[0].adj_hue - 25.0 / 60.0) < 1e-6 and abs(cube[0].adj_contrast - 1.4) < 1e-6
    # clouds / marble / wood: the constructors' members, the basis newNoise__ would build, and the lattice tables attached as data
    y = pc.EXACT["tex_dl_noise"]()
    d, keep = y.scene_desc()
    tex = [d.textures[i] for i in range(d.n_textures)]
    assert sorted(t.noise_type for t in tex if t.type >= capi.TEX_CLOUDS) == [0, 0, 1, 1, 2, 2, 2, 3, 4, 5, 6, 7, 8, 9]
    marble = [t for t in tex if t.type == capi.TEX_MARBLE and t.noise_type == capi.NOISE["blender"]][0]
    assert abs(marble.sharpness - 1.0 / 3.0) < 1e-7 and marble.noise_depth == 3 and marble.turbulence == 4.0 and marble.wave_shape == 0
    assert [t.sharpness for t in tex if t.type == capi.TEX_MARBLE and t.wave_shape == 2] == [1.0]   # "sharpness" <= 1 is ignored (texture_basic.cc:183)
    assert sorted(t.clouds_bias for t in tex if t.type == capi.TEX_CLOUDS and t.noise_type == 2) == [1, 2]
    assert sum(t.wood_rings for t in tex) == 1
    assert d.noise_tables_bytes == capi.NOISE_TABLES_BYTES == 13340
    blob = np.frombuffer(ctypes.string_at(d.noise_tables, d.noise_tables_bytes), np.uint8)
    hash_tab = blob[4 * (768 + 768 + 1542):][:512]
    assert sorted(hash_tab[:256].tolist()) == list(range(256)) and np.array_equal(hash_tab[:256], hash_tab[256:])   # a permutation, stored twice
    y = pc.EXACT["tex_dl_mask"]()
    d, keep = y.scene_desc()
    masks = [d.materials[i] for i in range(d.n_materials) if d.materials[i].type == capi.MAT_MASK]
    assert len(masks) == 3 and all(m.additional_depth == 0 and m.sampling_factor == 1.0 for m in masks)   # the factory reads neither
    from libyafaray_b200.interface import Interface
    z = Interface()
    with pytest.raises(NotImplementedError):
        z.addInstance(1, np.eye(4))


def C_string(ptr, n):
    return ctypes.string_at(ptr, n)


def test_golden_fixtures_present_and_sane():
    for name in pc.EXACT:
        f = np.load(os.path.join(ou.GOLDEN, name + ".npz"))["film"]
        assert f.ndim == 3 and f.shape[2] == 5 and np.isfinite(f).all() and f[..., 4].min() > 0


@pytest.mark.skipif(ou.ref_bin("ieee") is None, reason="oracle/_ref not built")
def test_oracle_reproduces_golden_film():
    """The committed fixture is what the reference CPU code renders (single thread, bit-reproducible)."""
    name = "simple_dl_gauss"
    film, _ = ou.run_reference_render(pc.EXACT[name]())
    gold = np.load(os.path.join(ou.GOLDEN, name + ".npz"))["film"]
    assert np.array_equal(film, gold)


@pytest.mark.skipif(ou.ref_bin("ieee") is None, reason="oracle/_ref not built")
def test_oracle_kat_matches_independent_restatement():
    with tempfile.TemporaryDirectory() as td:
        p = os.path.join(td, "kat.bin")
        subprocess.check_call([ou.ref_bin("ieee"), "kat", p], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        kat = ou.read_kat(p)
    for i in range(0, 4096, 37):
        x = (i * 2654435761 + i) & 0xFFFFFFFF
        h = qmc_ref.fnv32a(x)
        assert kat["fnv32"][i] == h
        assert kat["riVdC"][i] == qmc_ref.ri_vdc(x, h)
        assert kat["riS"][i] == qmc_ref.ri_s(x, h)
        assert kat["riLp"][i] == qmc_ref.ri_lp((x + h) & 0xFFFFFFFF)
        assert kat["riVdC0"][i] == qmc_ref.ri_vdc(i)
    # filter tables: box is all ones; gauss decreases from the centre
    ft = kat["filter_tables"].reshape(4, 16, 16)
    assert (ft[0] == 1).all() and ft[2][0, 0] > ft[2][8, 8] >= 0


def test_tile_owner_map_partitions_image():
    from libyafaray_b200 import parallel
    om = parallel.owner_map(100, 70, 32, 3)
    assert om.shape == (70, 100) and set(np.unique(om)) == {0, 1, 2}
    # blocks of 8 x 8 pixels, owner (tx + 2 ty) mod 3 (stride 2: nearest integer to 0.618 * 3 that is coprime with 3)
    assert om[0, 0] == 0 and om[0, 8] == 1 and om[0, 16] == 2 and om[0, 24] == 0 and om[8, 0] == 2 and om[7, 7] == 0
    assert [parallel.partition_stride(n) for n in (1, 2, 3, 4, 5, 6, 7, 8)] == [1, 1, 2, 3, 3, 5, 4, 5]
    counts = np.bincount(parallel.owner_map(1920, 1080, 32, 8).ravel())
    assert counts.max() - counts.min() <= 64 * 8   # every rank owns the same share of the frame up to a few blocks


def _gloo_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from libyafaray_b200 import parallel
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    w, h, ts = 96, 64, 32
    rng = np.random.default_rng(7)
    full = rng.random((h, w, 5)).astype(np.float32)
    own = parallel.owner_map(w, h, ts, world) == rank
    mine = full * own[..., None]
    # filter spill: every rank also deposits a little into its neighbours' pixels
    spill = np.roll(mine, 1, axis=1) * 0.25
    film = torch.from_numpy(mine + spill)
    parallel.reduce_film(film, 0)
    if rank == 0:
        want = full + np.roll(full, 1, axis=1) * 0.25
        q.put(float(np.abs(film.numpy() - want).max()))
    dist.destroy_process_group()


def test_film_reduce_world2_gloo():
    import torch.multiprocessing as mp
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    port = 29500 + (os.getpid() % 400)
    ps = [ctxm.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    for p in ps:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) < 1e-6


@pytest.mark.skipif(ou.gpu_frontend_bin() is None, reason="oracle/_ref/yaf_xml_gpu not built")
def test_gpu_frontend_fails_loudly_without_gpu_and_can_be_disabled():
    """The reference front end with GpuIntegrator linked in: without a GPU the render FAILS (no CPU fallback); with
    YAFGPU_DISABLE=1 the reference's own integrator is left in place and reproduces the golden film."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    y = pc.EXACT["simple_dl_gauss"]()
    film, info, r = ou.run_gpu_frontend(y)
    assert r.returncode != 0 and b"no CPU fallback" in r.stderr + r.stdout
    film, info, r = ou.run_gpu_frontend(y, env={"YAFGPU_DISABLE": "1"})
    assert r.returncode == 0
    gold = np.load(os.path.join(ou.GOLDEN, "simple_dl_gauss.npz"))["film"]
    assert np.array_equal(film, gold)


# ---- YAF_FILMv1: the reference's on-disk film format (imagefilm.cc:1340-1465,1560-1657) -------------------------------------
REF_FILM = os.path.join(ou.GOLDEN, "ref_cornell_32x18.film")


def test_film_file_reader_and_writer_against_a_film_saved_by_the_reference():
    """tests/golden/ref_cornell_32x18.film was written by the reference's own ImageFilm::imageFilmSave (tools/make_film_golden.py); the
    .npz beside it is the raw accumulator dump of the same render."""
    from libyafaray_b200 import film_io
    raw = np.load(os.path.join(ou.GOLDEN, "ref_cornell_32x18.npz"))["film"]
    f = film_io.read_film_file(REF_FILM, expect_shape=raw.shape[:2])
    assert (f.computer_node, f.base_sampling_offset, f.sampling_offset, f.cx0, f.cy0) == (3, 17, 2, 0, 0)
    assert len(f.passes) == 1 and len(f.aux_passes) == 0
    assert np.array_equal(f.film, raw)
    with tempfile.TemporaryDirectory() as td:
        p = film_io.write_film_file(os.path.join(td, "w.film"), raw, sampling_offset=2, base_sampling_offset=17, computer_node=3)
        assert open(p, "rb").read() == open(REF_FILM, "rb").read()          # byte for byte what the reference wrote
    assert film_io.film_path("/x/img.tga", 3) == "/x/img.tga - node 0003.film"


def test_film_file_reader_refuses_what_the_reference_refuses():
    from libyafaray_b200 import film_io
    data = open(REF_FILM, "rb").read()
    with tempfile.TemporaryDirectory() as td:
        bad = os.path.join(td, "bad.film")
        open(bad, "wb").write(b"YAF_FILMv2" + data[10:])
        with pytest.raises(ValueError):
            film_io.read_film_file(bad)                                      # header check (imagefilm.cc:1352-1358)
        open(bad, "wb").write(data[:-8])
        with pytest.raises(ValueError):
            film_io.read_film_file(bad)                                      # truncated pixel data
        with pytest.raises(ValueError):
            film_io.read_film_file(REF_FILM, expect_shape=(18, 33))          # "film check failed. Image width" (:1364-1369)


def test_films_of_a_folder_are_summed_like_imageFilmLoadAllInFolder():
    from libyafaray_b200 import film_io
    raw = np.load(os.path.join(ou.GOLDEN, "ref_cornell_32x18.npz"))["film"]
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "img.tga")
        film_io.write_film_file(film_io.film_path(out, 0), raw, sampling_offset=2, base_sampling_offset=5)
        film_io.write_film_file(film_io.film_path(out, 1), raw * np.float32(0.5), sampling_offset=9, base_sampling_offset=1, computer_node=1)
        film_io.write_film_file(os.path.join(td, "other - node 0000.film"), raw)            # other base name: ignored
        film_io.write_film_file(os.path.join(td, "img_small.film"), raw[:9])                # wrong size: skipped with a warning
        total, loaded = film_io.load_films_in_folder(out, raw.shape[:2])
        assert [os.path.basename(p) for p in loaded] == ["img.tga - node 0000.film", "img.tga - node 0001.film"]
        assert np.array_equal(total.film, raw + raw * np.float32(0.5))
        assert (total.sampling_offset, total.base_sampling_offset) == (9, 5)
        assert film_io.load_films_in_folder(os.path.join(td, "nothing.tga"), raw.shape[:2])[0] is None


@pytest.mark.skipif(ou.ref_bin() is None, reason="live oracle not available")
def test_reference_resumes_from_a_film_written_by_our_writer():
    """ImageFilm::imageFilmLoadAllInFolder of the reference reads a file produced by film_io.write_film_file and continues the render
    from it: same final film as when it resumes from the file it saved itself."""
    from libyafaray_b200 import film_io
    finals = []
    for ours in (False, True):
        with tempfile.TemporaryDirectory() as td:
            out = os.path.join(td, "img.tga")
            first, _ = ou.run_reference_render(pc.film_io_scene(1, "save"), film_path=out)
            fp = film_io.film_path(out, 3)
            if ours:
                f = film_io.read_film_file(fp)
                os.remove(fp)
                film_io.write_film_file(fp, first, f.sampling_offset, f.base_sampling_offset, f.computer_node)
            final, _ = ou.run_reference_render(pc.film_io_scene(3, "load-save"), film_path=out)
            saved = film_io.read_film_file(fp)
            assert np.array_equal(saved.film, final) and saved.sampling_offset == 2 + 3 + 3
            assert os.path.exists(fp + "-previous.bak")
            finals.append(final)
    assert np.array_equal(finals[0], finals[1])
    assert not np.array_equal(finals[0], first) and finals[0][..., 4].min() > first[..., 4].max() * 2


@pytest.mark.skipif(ou.ref_bin() is None, reason="oracle not built")
def test_mwc_restatement_matches_the_references_generator():
    """tests/ray_sets.py draws the scale ray sets from a restatement of Random (util_mcqmc.h:173-190): pinned on the KAT dump."""
    from tests import ray_sets as rs
    with tempfile.TemporaryDirectory() as td:
        p = os.path.join(td, "kat.bin")
        subprocess.check_call([ou.ref_bin(), "kat", p], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        k = ou.read_kat(p)["random"]
    for i, s in enumerate((0, 1, 123, (1804289383 + 123) & 0xffffffff)):
        assert np.array_equal(k[64 * i:64 * i + 64], rs.mwc_stream(s, 64))
        assert np.array_equal(k[64 * i:64 * i + 64], rs.mwc_reference_loop(s, 64))


@pytest.mark.skipif(ou.ref_bin() is None, reason="oracle not built")
def test_bulk_mesh_xml_export_reaches_the_references_loader():
    """addTriMeshArrays meshes are streamed to XML (to_xml) and load through the reference's unchanged mesh state
    (import_xml.cc:401-416,550-630): same triangle count, float32 coordinates survive the text round trip (the hits of a ray set
    equal those of the same mesh entered vertex by vertex), per-triangle material runs become <set_material> elements."""
    from libyafaray_b200 import scenes
    from libyafaray_b200.interface import Interface
    y = scenes.procedural_mesh(5000, 32, 18, 1, integrator="directlighting", mesh_lights=1, blend=True)
    rays = pc.trace_rays(4096, (-10, -10, 0.5), (10, 10, 8), seed=4)
    rays[:, 6] = -np.abs(rays[:, 6]) - 0.2
    with tempfile.TemporaryDirectory() as td:
        xml = y.to_xml(os.path.join(td, "s.xml"))
        txt = open(xml).read()
        assert txt.count("<set_material") > 10 and txt.count("<f a=") == sum(m.n_tris() for m in y.meshes.values())
        assert txt.count("<n x=") == len(y.meshes[1].np_points)
    h = ou.reference_hits(y, rays, 0)
    assert (h["tri"] >= 0).mean() > 0.3
    # the same geometry through startTriMesh/addVertex/addTriangle: identical hits
    y2 = Interface()
    y2.materials, y2.lights, y2.cameras, y2.backgrounds, y2.integrators = y.materials, y.lights, y.cameras, y.backgrounds, y.integrators
    y2.render_params = y.render_params
    y2.startGeometry()
    for oid in sorted(y.meshes):
        m = y.meshes[oid]
        if m.np_faces is None:
            y2.meshes[oid] = m
            continue
        names, per_tri = ([m.np_mat], np.zeros(len(m.np_faces), np.int32)) if isinstance(m.np_mat, str) else m.np_mat
        y2.startTriMesh(oid, len(m.np_points), len(m.np_faces), False, False, 0)
        for k, p in enumerate(m.np_points.tolist()):
            y2.addVertex(*p)
            if m.np_normals is not None:
                y2.addNormal(*m.np_normals[k].tolist())
        for f, t in zip(m.np_faces.tolist(), per_tri):
            y2.addTriangle(f[0], f[1], f[2], names[int(t)])
        y2.endTriMesh()
    y2.endGeometry()
    h2 = ou.reference_hits(y2, rays, 0)
    assert np.array_equal(h.view(np.uint8), h2.view(np.uint8))


def test_scale_golden_subsets_are_consistent_with_the_ray_generator():
    """The committed subset of the reference's answers at 10^6 triangles belongs to the ray set the generator produces today."""
    from tests import ray_sets as rs
    z = np.load(os.path.join(ou.GOLDEN, "scale_hits_c4.npz"))
    y = rs.scale_scene("c4")
    rays = rs.scale_rays("c4", y)
    assert len(rays) == 2 << 20 and np.isfinite(rays).all()
    assert bytes(z["digest"]).decode() == rs.rays_digest(rays)
    h = z["hits"].view(ou.HIT_DTYPE).reshape(-1)
    assert len(h) == len(z["index"]) and 0.3 < (h["tri"] >= 0).mean() < 0.95


def test_triangle_intersect_restatement_reproduces_the_references_hits():
    """tests/ray_sets.py triangle_intersect_f32 (the certificate behind the epsilon-tie rule of the scale tests) returns the reference's
    own t,u,v bit for bit on every hit of the committed 10^6-triangle subset."""
    from tests import ray_sets as rs
    z = np.load(os.path.join(ou.GOLDEN, "scale_hits_c4.npz"))
    y = rs.scale_scene("c4")
    rays = rs.scale_rays("c4", y)[z["index"]]
    h = z["hits"].view(ou.HIT_DTYPE).reshape(-1)
    m = h["tri"] >= 0
    ok, t, u, v = rs.triangle_intersect_f32(*rs.scene_triangles(y), rays[m], h["tri"][m])
    assert ok.all() and m.sum() > 5000
    for a, b in ((t, "t"), (u, "u"), (v, "v")):
        assert np.array_equal(a.view(np.uint32), h[b][m].view(np.uint32)), b
