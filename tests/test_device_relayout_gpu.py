"""srt_upload_scene's re-layout done by kernels (csrc/srt_relayout.cu, SRT_OPT_DEVICE_RELAYOUT) against the host walk.

The two produce the same records up to where a node pair sits in its array, so every render must come out with the same
bits, the same primary-hit ids and the same counters -- and both must equal the oracle.  Scenes: the reference's Rubik
asset, two models behind frames, a textured model, a soup above the host path's threading threshold, a soup big enough
that the default (2 = automatic) picks the device path by itself, a node array in another numbering (must fall back), and
invalid scenes (must be refused in the host walk's words, the previous scene staying in place).
"""
import numpy as np
import pytest

import scenes
from test_parity_gpu import assert_bit_exact, cpu_render, framed, gpu_render

pytestmark = pytest.mark.gpu


def both_ways(pkg, ctx, scene, W, H, frames, expect_device=True, **kw):
    out = []
    for mode in (0, 1):
        ctx.set_option(pkg.OPT_DEVICE_RELAYOUT, mode)
        acc, img, ids, cnt = gpu_render(pkg, ctx, W, H, True, frames, pkg.TRAV_REFERENCE, scene=scene, **kw)
        on_device, ms = ctx.upload_info()
        assert on_device == (mode == 1 and expect_device)
        out.append((acc, img, ids, cnt))
    ctx.set_option(pkg.OPT_DEVICE_RELAYOUT, 2)
    (a0, i0, d0, c0), (a1, i1, d1, c1) = out
    assert np.array_equal(d0, d1)
    assert_bit_exact(a1, a0)
    assert np.array_equal(i0, i1)
    assert c0["segments"] == c1["segments"] and c0["shadow_rays"] == c1["shadow_rays"]
    return out[1]


@pytest.mark.parametrize("schedule", [3, 30, 31])
def test_rubik_device_relayout_equals_host_and_oracle(pkg, ctx, oracle, schedule):
    scene = scenes.rubik_scene()
    acc, img, ids, cnt = both_ways(pkg, ctx, scene, 320, 200, 3, schedule=schedule)
    oacc, oimg, oids, ocnt = cpu_render(oracle, 320, 200, True, 3, scene=scene)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]


def test_two_models_with_frames(pkg, ctx, oracle, tmp_path):
    name, _ = scenes.write_airplane_standin(str(tmp_path), detail=24)
    scene = scenes.concat_scenes(scenes.rubik_scene(), scenes.load_obj_scene(str(tmp_path), name, str(tmp_path)))
    scene["bvhs"]["frame"][0] = framed(0.6, 30.0, (-7.0, 2.0, 0.0))
    scene["bvhs"]["frame"][1] = framed(1.3, -20.0, (6.0, 1.0, 5.0))
    acc, img, ids, cnt = both_ways(pkg, ctx, scene, 256, 128, 2)
    oacc, _, oids, _ = cpu_render(oracle, 256, 128, True, 2, scene=scene)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    # the frame of a model is replaced in the records the device built, like in the host's
    moved = framed(1.3, 75.0, (2.0, 4.0, -3.0))
    ctx.update_frame(1, moved)
    ctx.resize(256, 128)
    ctx.enable_primary_ids(True)
    ctx.set_params(pkg.make_params(256, 128, scenes.reset_camera(True), True, accum_frames=1, bvh_count=2))
    ctx.render(2)
    scene["bvhs"]["frame"][1] = moved
    oacc2, _, oids2, _ = cpu_render(oracle, 256, 128, True, 2, scene=scene)
    assert np.array_equal(ctx.read_primary_ids(), oids2)
    assert_bit_exact(ctx.read_accum(), oacc2)


@pytest.mark.parametrize("schedule", [3, 30])
def test_soup_device_relayout(pkg, ctx, oracle, tmp_path, schedule):
    scene, info = scenes.soup_scene(70000, str(tmp_path))
    cam = scenes.reset_camera(True, origin=(0.0, 0.0, 120.0))
    acc, img, ids, cnt = both_ways(pkg, ctx, scene, 192, 104, 2, cam=cam, max_depth=8, schedule=schedule)
    oacc, _, oids, ocnt = cpu_render(oracle, 192, 104, True, 2, scene=scene, cam=cam, max_depth=8)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert cnt["segments"] == ocnt["segments"]


def test_automatic_choice_above_the_finder_limit(pkg, ctx, oracle, tmp_path):
    """Default option value 2: a scene of more than 2 Mi triangles is re-laid out on the device without being asked; its
    picture equals the host walk's and the oracle's (configs[4]'s kind of scene at a fifth of its size: 2.2 M triangles,
    4.4 M nodes, maxDepth 8)."""
    scene, info = scenes.soup_scene((1 << 21) + 100000, str(tmp_path))
    cam = scenes.reset_camera(True, origin=(0.0, 0.0, 120.0))
    ctx.set_option(pkg.OPT_DEVICE_RELAYOUT, 2)
    a1, i1, d1, c1 = gpu_render(pkg, ctx, 160, 88, True, 1, pkg.TRAV_REFERENCE, scene=scene, cam=cam, max_depth=8)
    on_device, ms = ctx.upload_info()
    assert on_device and ms > 0.0
    ctx.set_option(pkg.OPT_DEVICE_RELAYOUT, 0)
    a0, i0, d0, c0 = gpu_render(pkg, ctx, 160, 88, True, 1, pkg.TRAV_REFERENCE, scene=scene, cam=cam, max_depth=8)
    assert not ctx.upload_info()[0]
    ctx.set_option(pkg.OPT_DEVICE_RELAYOUT, 2)
    assert np.array_equal(d0, d1)
    assert_bit_exact(a1, a0)
    assert c0["segments"] == c1["segments"] and c0["shadow_rays"] == c1["shadow_rays"]
    oacc, _, oids, ocnt = cpu_render(oracle, 160, 88, True, 1, scene=scene, cam=cam, max_depth=8)
    assert np.array_equal(d1, oids)
    assert_bit_exact(a1, oacc)
    assert c1["segments"] == ocnt["segments"] and c1["shadow_rays"] == ocnt["shadow_rays"]


def reversed_numbering(scene):
    """The same tree with the node array written back to front in sibling pairs: children sit BELOW their parent, which
    is not the reference's numbering.  Root (old 0) moves to the end."""
    nodes = scene["nodes"]
    n = len(nodes)
    assert n % 2 == 1  # root + sibling pairs
    new_of_old = np.empty(n, np.int64)
    new_of_old[0] = n - 1
    pairs = (n - 1) // 2
    for p in range(pairs):  # old pair (1+2p, 2+2p) -> new pair at the mirrored position, order inside the pair kept
        q = pairs - 1 - p
        new_of_old[1 + 2 * p] = 2 * q
        new_of_old[2 + 2 * p] = 2 * q + 1
    out = nodes.copy()
    out[new_of_old] = nodes
    inner = out["count"] == 0
    out["index"][inner] = new_of_old[out["index"][inner]]
    s = dict(scene)
    s["nodes"] = np.ascontiguousarray(out)
    s["bvhs"] = scene["bvhs"].copy()
    s["bvhs"]["first_index"][0] = n - 1
    return s


def test_other_numbering_falls_back_to_the_host_walk(pkg, ctx, oracle):
    scene = scenes.rubik_scene()
    assert len(scene["bvhs"]) == 1
    other = reversed_numbering(scene)
    acc, img, ids, cnt = both_ways(pkg, ctx, other, 200, 120, 2, expect_device=False)
    oacc, _, oids, _ = cpu_render(oracle, 200, 120, True, 2, scene=scene)  # the same tree: the same picture
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)


def test_textured_model_device_relayout(pkg, ctx, tmp_path):
    """The textured-albedo option needs, per triangle, the BVH whose frame gives the model-space hit point: computed by the
    device sweep too."""
    import test_texture_option as tt
    scene, textures = tt.textured_scene(str(tmp_path))
    ctx.set_option(pkg.OPT_TEXTURES, 1)
    try:
        both_ways(pkg, ctx, scene, 200, 120, 2, textures=textures)
    finally:
        ctx.set_option(pkg.OPT_TEXTURES, 0)


def test_invalid_scene_refused_in_the_host_walks_words_and_previous_scene_kept(pkg, ctx, tmp_path):
    good = scenes.rubik_scene()
    ctx.set_option(pkg.OPT_DEVICE_RELAYOUT, 1)
    try:
        ref = gpu_render(pkg, ctx, 160, 96, True, 1, pkg.TRAV_REFERENCE, scene=good)
        soup, _ = scenes.soup_scene(70000, str(tmp_path))
        bad = {k: v.copy() for k, v in soup.items()}
        bad["triangles"]["v"][1234, 2] = len(soup["vertices"]) + 5
        bad["triangles"]["material"][77] = len(soup["materials"]) + 1
        with pytest.raises(pkg.SrtError, match="triangle 77: material index out of range"):
            ctx.upload_scene(bad)
        bad["triangles"]["material"][77] = 0
        with pytest.raises(pkg.SrtError, match="triangle 1234: vertex index out of range"):
            ctx.upload_scene(bad)
        bad = {k: v.copy() for k, v in soup.items()}
        leaf = np.nonzero(bad["nodes"]["count"] > 0)[0][5]
        bad["nodes"]["index"][leaf] = len(soup["triangles"])
        with pytest.raises(pkg.SrtError, match="out of range"):
            ctx.upload_scene(bad)
        # the Rubik scene is still the one in place
        ctx.resize(160, 96)  # fresh accumulation image
        ctx.enable_primary_ids(True)
        ctx.set_params(pkg.make_params(160, 96, scenes.reset_camera(True), True, accum_frames=1, bvh_count=1))
        ctx.render(1)
        assert np.array_equal(ref[2], ctx.read_primary_ids())
        assert_bit_exact(ctx.read_accum(), ref[0])
    finally:
        ctx.set_option(pkg.OPT_DEVICE_RELAYOUT, 2)
