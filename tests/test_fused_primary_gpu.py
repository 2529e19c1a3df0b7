"""SRT_OPT_FUSED_PRIMARY (off by default: measured slower, include/srt_b200.h): where every path of a launch starts in its first wave the hit finder's find kernel makes the
camera rays itself (no wf_init_kernel), gives a slot only to a ray that comes back with candidate triangles and ends any
other with one bit of the culled map (srt_wf_find.cuh, PRIMARY).  Which rays get slots, and in which order, differs from the
init kernel's -- no result may: accumulation image, output, primary ids and the counters must be those of the option's
other value and of the oracle, bit for bit.  The launches here are forced through waves (SRT_OPT_TAIL_PATHS below the
default: at these sizes the default hands a launch to the tail kernel straight after the init kernel, which the option
leaves alone).
"""
import numpy as np
import pytest

import scenes
from test_parity_gpu import assert_bit_exact, cpu_render, framed, gpu_render

pytestmark = pytest.mark.gpu


def both(pkg, ctx, scene, W, H, frames, tail, **kw):
    out = []
    for fused in (0, 1):
        ctx.set_option(pkg.OPT_FUSED_PRIMARY, fused)
        out.append(gpu_render(pkg, ctx, W, H, True, frames, pkg.TRAV_REFERENCE, scene=scene, schedule=3, tail=tail, finder=1, **kw))
        assert ctx.finder_info()[0]
    ctx.set_option(pkg.OPT_FUSED_PRIMARY, 0)
    (a0, i0, d0, c0), (a1, i1, d1, c1) = out
    assert np.array_equal(d0, d1)
    assert_bit_exact(a1, a0)
    assert np.array_equal(i0, i1)
    for key in ("segments", "shadow_rays", "paths"):
        assert c0[key] == c1[key], key
    return out[1]


@pytest.mark.parametrize("tail", [0, 5000])
def test_rubik(pkg, ctx, oracle, tail):
    scene = scenes.rubik_scene()
    acc, img, ids, cnt = both(pkg, ctx, scene, 320, 200, 3, tail)
    oacc, oimg, oids, ocnt = cpu_render(oracle, 320, 200, True, 3, scene=scene)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert np.array_equal(img, oimg)
    assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]


def test_later_frames_and_a_band_of_rows(pkg, ctx, oracle):
    """accum_frames > 1 (the noise offset of GetRay) and a context that renders rows 40..136 only (a rank of a band split)."""
    scene = scenes.rubik_scene()
    acc, img, ids, cnt = both(pkg, ctx, scene, 256, 144, 2, 0, first_frame=5, rows=(40, 136))
    ctx.set_rows(0, 144)
    oacc, _, oids, _ = cpu_render(oracle, 256, 144, True, 2, scene=scene, first_frame=5)
    assert np.array_equal(ids[40:136], oids[40:136])
    assert_bit_exact(acc[40:136], oacc[40:136])


def test_camera_inside_the_model(pkg, ctx, oracle):
    """No camera ray is an outside ray: every one collects candidates without the distance cull."""
    scene = scenes.rubik_scene()
    cam = scenes.reset_camera(True, origin=(0.1, 0.2, 0.3))
    acc, img, ids, cnt = both(pkg, ctx, scene, 192, 104, 2, 0, cam=cam)
    oacc, _, oids, ocnt = cpu_render(oracle, 192, 104, True, 2, scene=scene, cam=cam)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert cnt["segments"] == ocnt["segments"]


def test_two_models(pkg, ctx, oracle):
    """Two BVHs: a camera ray enters the second tree after the first, its ray made again from the path number."""
    scene = scenes.concat_scenes(scenes.rubik_scene(), scenes.rubik_scene())
    scene["bvhs"]["frame"][0] = framed(0.6, 30.0, (-5.0, 1.0, 0.0))
    scene["bvhs"]["frame"][1] = framed(0.5, -20.0, (4.0, -1.0, 2.0))
    acc, img, ids, cnt = both(pkg, ctx, scene, 256, 128, 2, 0)
    oacc, _, oids, ocnt = cpu_render(oracle, 256, 128, True, 2, scene=scene)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert cnt["segments"] == ocnt["segments"]


def test_standin_with_big_leaves(pkg, ctx, oracle, tmp_path):
    """Leaves of up to 165 triangles: camera rays with more candidates than a lane keeps go to the overflow walk with the
    slot they were given a moment before."""
    name, _ = scenes.write_airplane_standin(str(tmp_path), detail=48)
    scene = scenes.load_obj_scene(str(tmp_path), name, str(tmp_path))
    acc, img, ids, cnt = both(pkg, ctx, scene, 256, 144, 2, 0)
    oacc, _, oids, ocnt = cpu_render(oracle, 256, 144, True, 2, scene=scene)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert cnt["segments"] == ocnt["segments"]


def test_without_ids_on_two_streams(pkg, ctx, oracle):
    """The library's default way: no primary ids, SRT_OPT_OVERLAP on, calls queued back to back, one-frame chunks."""
    W, H = 256, 144
    scene = scenes.rubik_scene()
    unit, uniform = scenes.noise_tables(W, H)
    results = []
    for fused in (0, 1):
        ctx.set_option(pkg.OPT_FUSED_PRIMARY, fused)
        ctx.set_option(pkg.OPT_OVERLAP, 1)
        ctx.set_option(pkg.OPT_TRAVERSAL, pkg.TRAV_REFERENCE)
        ctx.set_option(pkg.OPT_SCHEDULE, 3)
        ctx.set_option(pkg.OPT_FINDER, 1)
        ctx.set_option(pkg.OPT_TAIL_PATHS, 3000)
        ctx.set_option(pkg.OPT_SAMPLE_BUDGET_MB, 1)
        ctx.upload_scene(scene)
        ctx.upload_noise(unit, uniform)
        ctx.set_lights(scenes.model_lights())
        ctx.resize(W, H)
        ctx.enable_primary_ids(False)
        ctx.set_params(pkg.make_params(W, H, scenes.reset_camera(True), True, accum_frames=1, bvh_count=1))
        ctx.render_async(3)
        ctx.render_async(2)
        ctx.sync()
        results.append((ctx.read_accum(), ctx.read_output()))
    ctx.set_option(pkg.OPT_FUSED_PRIMARY, 0)
    assert_bit_exact(results[1][0], results[0][0])
    assert np.array_equal(results[1][1], results[0][1])
    oacc, oimg, _, _ = cpu_render(oracle, W, H, True, 5, scene=scene)
    assert_bit_exact(results[1][0], oacc)
    assert np.array_equal(results[1][1], oimg)
