"""SRT_OPT_OVERLAP: a chunk's waves on an internal stream with one of two sets of path slots, its tail kernel and resolve on
the context's stream, so that the tail of one chunk (or of one srt_render call) runs beside the waves of the next.  The
accumulation image is only ever touched on the context's stream, in call order: every picture must have the bits of the
plain one-stream schedule and of the oracle -- with several chunks per call, with calls queued back to back without a
synchronisation in between, with and without resets, with uploads in between.  (Primary ids switch the overlap off, so
nothing here records them.)
"""
import numpy as np
import pytest

import scenes
from test_parity_gpu import assert_bit_exact, cpu_render

pytestmark = pytest.mark.gpu

W, H = 256, 144


def setup(pkg, ctx, scene, overlap, tail, wave_slots=None, budget_mb=None):
    ctx.set_option(pkg.OPT_OVERLAP, overlap)
    ctx.set_option(pkg.OPT_TRAVERSAL, pkg.TRAV_REFERENCE)
    ctx.set_option(pkg.OPT_SCHEDULE, 3)
    ctx.set_option(pkg.OPT_TAIL_PATHS, tail)
    if wave_slots:
        ctx.set_option(pkg.OPT_WAVE_SLOTS, wave_slots)
    if budget_mb:
        ctx.set_option(pkg.OPT_SAMPLE_BUDGET_MB, budget_mb)
    unit, uniform = scenes.noise_tables(W, H)
    ctx.upload_scene(scene)
    ctx.upload_noise(unit, uniform)
    ctx.set_lights(scenes.model_lights())
    ctx.resize(W, H)


def params(pkg, scene, first, reset=False):
    return pkg.make_params(W, H, scenes.reset_camera(True), True, accum_frames=first, bvh_count=len(scene["bvhs"]),
                           reset=1 if reset else 0)


@pytest.mark.parametrize("tail", [0, 3000, 1 << 20])
def test_chunks_and_calls_back_to_back(pkg, ctx, oracle, tail):
    """6 frames as 3 + 2 + 1 in three calls queued without a sync, each call cut into one-frame chunks (1 MB sample budget:
    36 864 paths x 16 B): tails and waves of neighbouring chunks and calls overlap."""
    scene = scenes.rubik_scene()
    results = []
    for overlap in (0, 1):
        setup(pkg, ctx, scene, overlap, tail, budget_mb=1)
        ctx.set_params(params(pkg, scene, 1))
        ctx.render_async(3)
        ctx.render_async(2)
        ctx.render_async(1)
        ctx.sync()
        results.append((ctx.read_accum(), ctx.read_output(), ctx.counters()))
    (a0, o0, c0), (a1, o1, c1) = results
    assert_bit_exact(a1, a0)
    assert np.array_equal(o0, o1)
    oacc, oout, _, ocnt = cpu_render(oracle, W, H, True, 6, scene=scene)
    assert_bit_exact(a1, oacc)
    assert np.array_equal(o1, oout)


def test_small_pool_regeneration_and_reset_between_calls(pkg, ctx, oracle):
    """A pool smaller than a chunk (slots take new paths as old ones end), a reset dispatch between two calls: the clear
    must come after the first call's resolve and before the second one's."""
    scene = scenes.rubik_scene()
    results = []
    for overlap in (0, 1):
        setup(pkg, ctx, scene, overlap, 2000, wave_slots=8192)
        ctx.set_params(params(pkg, scene, 1))
        ctx.render_async(2)
        ctx.set_params(params(pkg, scene, 1, reset=True))
        ctx.render_async(1)  # the reset dispatch: clears, renders nothing
        ctx.set_params(params(pkg, scene, 1))
        ctx.render_async(3)
        ctx.sync()
        results.append(ctx.read_accum())
    assert_bit_exact(results[1], results[0])
    oacc, _, _, _ = cpu_render(oracle, W, H, True, 3, scene=scene)
    assert_bit_exact(results[1], oacc)


def test_upload_between_calls(pkg, ctx, oracle, tmp_path):
    """Another scene between two calls: the second call's waves must see it, the first call's tail must not."""
    rubik = scenes.rubik_scene()
    soup, _ = scenes.soup_scene(5000, str(tmp_path))
    cam = scenes.reset_camera(True)
    setup(pkg, ctx, rubik, 1, 3000, budget_mb=1)
    ctx.set_params(params(pkg, rubik, 1))
    ctx.render_async(2)
    first = ctx.read_accum()
    ctx.upload_scene(soup)
    ctx.set_params(params(pkg, soup, 1, reset=True))
    ctx.render_async(1)
    ctx.set_params(params(pkg, soup, 1))
    ctx.render_async(2)
    second = ctx.read_accum()
    o1, _, _, _ = cpu_render(oracle, W, H, True, 2, scene=rubik, cam=cam)
    o2, _, _, _ = cpu_render(oracle, W, H, True, 2, scene=soup, cam=cam)
    assert_bit_exact(first, o1)
    assert_bit_exact(second, o2)
