"""Multi-GPU behind the boundary (`-m gpu`): yafgpu_group — replicated scene, interleaved-tile partition, film sum on the first device —
through the C ABI and through the reference front end (yaf_xml_gpu with YAFGPU_DEVICES).

Film-sum semantics: ImageFilm::imageFilmLoadAllInFolder (/root/reference/src/common/imagefilm.cc:1506-1535), tile grid: ImageSplitter
(/root/reference/src/common/imagesplitter.cc:37-50).  The N-GPU film must equal the 1-GPU film up to float summation order (<= 1e-4 of
the largest accumulator; filter weights exactly equal for the box filter).  On a single-GPU box the same path runs with several group
members sharing the device (YAFGPU_FLAG_SHARE_DEVICE); the tests that need two physical GPUs skip there."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import oracle_util as ou, parity_cases as pc  # noqa: E402

pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch
    return torch.cuda.device_count()


def _same_film(a, b):
    assert a.shape == b.shape
    assert np.array_equal(a[..., 4], b[..., 4])                                   # box filter: integer weights
    assert np.abs(a[..., :4] - b[..., :4]).max() <= 1e-4 * float(np.abs(a[..., :4]).max())


def _scene():
    from libyafaray_b200 import scenes
    return scenes.cornell_box(200, 120, 8, path_samples=1, bounces=4, raydepth=4, aa_passes=2, inc_samples=4)


def test_group_of_one_is_a_plain_context():
    from libyafaray_b200 import Context, Group, render_scene
    y = _scene()
    c, g = Context(0), Group([0])
    try:
        assert g.size() == 1
        _same_film(render_scene(y, ctx=g), render_scene(y, ctx=c))
    finally:
        c.close(); g.close()


@pytest.mark.parametrize("members", [2, 3])
def test_group_members_sharing_one_gpu_sum_to_the_single_context_film(members):
    """Partition by interleaved tiles + film sum, exercised on ONE device (several contexts on it): two AA passes, 200x120 (tiles do not
    divide the frame), path tracing."""
    from libyafaray_b200 import Context, Group, capi, render_scene
    y = _scene()
    c = Context(0)
    g = Group([0] * members, capi.FLAG_SHARE_DEVICE)
    try:
        assert g.size() == members
        ref = render_scene(y, ctx=c)
        got = render_scene(y, ctx=g)
        _same_film(got, ref)
        st = g.stats()
        assert st["film_reduces"] >= 2 and st["camera_samples"] == c.stats()["camera_samples"]
        # a second frame through the same group: the previous reduction must not leak into it
        _same_film(render_scene(y, ctx=g), ref)
    finally:
        c.close(); g.close()


def test_group_decides_adaptive_passes_on_the_summed_film():
    """yafgpu_group_next_pass: the resampling mask comes from the SUMMED film on the first device and every member renders its tiles of it; the result is the
    reference's adaptive film (tests/golden/cornell_dl_adaptive_variance.npz: four passes, variance windows, resampled floor)."""
    from libyafaray_b200 import Group, capi, render_scene
    from tests.test_gpu import _assert_film_matches
    name = "cornell_dl_adaptive_variance"
    gold = np.load(os.path.join(ou.GOLDEN, name + ".npz"))["film"]
    g = Group([0, 0, 0], capi.FLAG_SHARE_DEVICE)
    try:
        _assert_film_matches(render_scene(pc.ADAPTIVE[name](), ctx=g), gold)
    finally:
        g.close()


@pytest.mark.parametrize("parts", [2, 3, 8])
def test_device_partition_is_the_owner_map_of_the_python_mirror(parts):
    """part_index / part_count on the device (csrc/integrator.cu buildPixelList: 8 x 8 blocks, shifted rows) against libyafaray_b200/parallel.py owner_map,
    pixel by pixel: with a one-pixel box filter a sample only touches its own pixel, so the pixels that carry weight after rendering part k are exactly k's."""
    from libyafaray_b200 import Context, parallel, scenes
    w, h = 100, 70
    y = scenes.simple_point_scene(w, h, 1, filter_type="box", AA_pixelwidth=1.0)
    desc, keep = y.scene_desc()
    om = parallel.owner_map(w, h, desc.film.tile_size, parts)
    c = Context(0)
    try:
        c.upload(desc, keep)
        seen = np.zeros((h, w), np.int32)
        for k in range(parts):
            c.film_clear(); c.render_pass(1, 0, part_index=k, part_count=parts)
            mine = c.read_film()[..., 4] > 0
            assert np.array_equal(mine, om == k), (k, int((mine != (om == k)).sum()))
            seen += mine
        assert (seen == 1).all()
    finally:
        c.close()


def test_group_rejects_bad_device_lists():
    from libyafaray_b200 import Group, capi
    with pytest.raises(capi.YafGpuError, match="listed twice"):
        Group([0, 0])
    with pytest.raises(capi.YafGpuError, match="out of range"):
        Group([0, 99])


@pytest.mark.skipif(_n_gpus() < 2, reason="needs two GPUs")
def test_two_gpu_group_equals_single_gpu_film():
    from libyafaray_b200 import Context, Group, render_scene
    y = _scene()
    c, g = Context(0), Group([0, 1])
    try:
        _same_film(render_scene(y, ctx=g), render_scene(y, ctx=c))
    finally:
        c.close(); g.close()


@pytest.mark.skipif(ou.gpu_frontend_bin() is None, reason="yaf_xml_gpu not built")
def test_reference_frontend_on_a_device_group():
    """XML -> reference loader -> GpuIntegrator on YAFGPU_DEVICES -> reference ImageFilm: equal to the one-GPU run of the same binary.
    Two physical GPUs when the box has them, else the all-GPUs list of a one-GPU box (a group of one)."""
    y = pc.EXACT["cornell_dl_glass"]()
    one, _, r1 = ou.run_gpu_frontend(y, env={"YAFGPU_DEVICE": "0"})
    assert r1.returncode == 0, r1.stderr.decode(errors="replace")[-2000:]
    devs = "0,1" if _n_gpus() >= 2 else "all"
    many, _, r2 = ou.run_gpu_frontend(y, env={"YAFGPU_DEVICES": devs, "YAF_REF_VERBOSE": "1"})
    assert r2.returncode == 0, r2.stderr.decode(errors="replace")[-2000:]
    assert ("on %d GPU(s)" % (2 if _n_gpus() >= 2 else 1)) in (r2.stdout + r2.stderr).decode(errors="replace")
    assert np.abs(many[..., 4] - one[..., 4]).max() <= 1e-5 * one[..., 4].max()
    assert np.abs(many[..., :4] - one[..., :4]).max() <= 1e-4 * float(np.abs(one[..., :4]).max())
