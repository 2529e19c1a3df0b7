"""The hit finder's trees built by kernels (csrc/srt_finder_build.cu, SRT_OPT_DEVICE_FINDER_BUILD = 1, the default for
scenes of 16 Ki triangles and more; the environment variable SRT_FINDER_DEVICE_ALWAYS=1 set below lifts that limit so that
the small scenes here go through the kernels too) against the host builder (0).  The two trees are not the same tree (the
host builder samples big nodes), and they need not be: the finder collects candidate triangles, the replay of the shader's own culls decides.  So every picture must come
out with the same bits either way -- and both must equal the oracle -- with the finder forced on (schedule variant 31:
waves to the very end through the finder) and with the library's defaults.
"""
import numpy as np
import pytest

import scenes
from test_parity_gpu import assert_bit_exact, cpu_render, framed, gpu_render

pytestmark = pytest.mark.gpu



@pytest.fixture(autouse=True)
def device_builder_for_small_scenes(monkeypatch):
    monkeypatch.setenv("SRT_FINDER_DEVICE_ALWAYS", "1")  # read by srt_upload_scene at every call


def both_builders(pkg, ctx, scene, W, H, frames, schedule, **kw):
    out = []
    for mode in (0, 1):
        ctx.set_option(pkg.OPT_DEVICE_FINDER_BUILD, mode)
        out.append(gpu_render(pkg, ctx, W, H, True, frames, pkg.TRAV_REFERENCE, scene=scene, schedule=schedule, **kw))
        if schedule == 31:
            assert ctx.finder_info()[0]  # the finder answered the rays
    ctx.set_option(pkg.OPT_DEVICE_FINDER_BUILD, 1)
    (a0, i0, d0, c0), (a1, i1, d1, c1) = out
    assert np.array_equal(d0, d1)
    assert_bit_exact(a1, a0)
    assert np.array_equal(i0, i1)
    assert c0["segments"] == c1["segments"] and c0["shadow_rays"] == c1["shadow_rays"]
    return out[1]


@pytest.mark.parametrize("schedule", [3, 31])
def test_rubik(pkg, ctx, oracle, schedule):
    scene = scenes.rubik_scene()
    acc, img, ids, cnt = both_builders(pkg, ctx, scene, 320, 200, 3, schedule)
    oacc, _, oids, ocnt = cpu_render(oracle, 320, 200, True, 3, scene=scene)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]


@pytest.mark.parametrize("schedule", [3, 31])
def test_airplane_standin(pkg, ctx, oracle, tmp_path, schedule):
    name, _ = scenes.write_airplane_standin(str(tmp_path), detail=48)
    scene = scenes.load_obj_scene(str(tmp_path), name, str(tmp_path))
    acc, img, ids, cnt = both_builders(pkg, ctx, scene, 256, 144, 2, schedule)
    oacc, _, oids, ocnt = cpu_render(oracle, 256, 144, True, 2, scene=scene)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert cnt["segments"] == ocnt["segments"]


def test_two_models_and_a_tiny_one(pkg, ctx, oracle, tmp_path):
    """Two trees in one build, one of them the 12-triangle box of the adversarial scenes' kind (a root with few levels)."""
    from test_adversarial_gpu import box, write_obj
    verts, faces = [], []
    box(verts, faces, (-1.0, -1.0, -1.0), (1.5, 2.0, 1.0))
    name = write_obj(str(tmp_path), "tinybox", verts, [("m", 20.0, (0.7, 0.2, 0.1), (0.2, 0.2, 0.2), faces)])
    tiny = scenes.load_obj_scene(str(tmp_path), name, str(tmp_path))
    scene = scenes.concat_scenes(scenes.rubik_scene(), tiny)
    scene["bvhs"]["frame"][0] = framed(0.6, 30.0, (-7.0, 2.0, 0.0))
    scene["bvhs"]["frame"][1] = framed(0.4, -20.0, (3.0, 1.0, 2.0))
    acc, img, ids, cnt = both_builders(pkg, ctx, scene, 256, 128, 2, 31)
    oacc, _, oids, _ = cpu_render(oracle, 256, 128, True, 2, scene=scene)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)


def test_soup_with_the_finder_forced(pkg, ctx, oracle, tmp_path):
    scene, info = scenes.soup_scene(20000, str(tmp_path))
    cam = scenes.reset_camera(True, origin=(0.0, 0.0, 120.0))
    acc, img, ids, cnt = both_builders(pkg, ctx, scene, 192, 104, 2, 31, cam=cam, max_depth=8)
    oacc, _, oids, ocnt = cpu_render(oracle, 192, 104, True, 2, scene=scene, cam=cam, max_depth=8)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert cnt["segments"] == ocnt["segments"]
