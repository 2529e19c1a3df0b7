"""The textured-albedo option (include/srt_texmath.h): known answers for the shared arithmetic on the
CPU, and kernel-vs-oracle parity on the GPU.  The reference's own textured branch is undefined as shipped (decree Q7
keeps albedo = Kd by default), so there is no reference output to compare with: the option is pinned against an
independent numpy restatement of "barycentric texture coordinate + GL REPEAT/LINEAR sample of the base level"."""
import numpy as np
import pytest

import scenes


def numpy_bilinear(img, s, t):
    H, W = img.shape[:2]
    x = np.float64(s) * W - 0.5
    y = np.float64(t) * H - 0.5
    x0, y0 = np.floor(x), np.floor(y)
    fx, fy = x - x0, y - y0
    x0, y0 = int(x0) % W, int(y0) % H
    x1, y1 = (x0 + 1) % W, (y0 + 1) % H
    c = lambda yy, xx: img[yy, xx, :3].astype(np.float64) / 255.0
    return (c(y0, x0) * (1 - fx) + c(y0, x1) * fx) * (1 - fy) + (c(y1, x0) * (1 - fx) + c(y1, x1) * fx) * fy


def test_bilinear_known_answers(oracle):
    rng = np.random.default_rng(7)
    img = rng.integers(0, 256, (5, 7, 4), dtype=np.uint8)
    H, W = img.shape[:2]
    # texel centres return the texel itself, exactly; one period away is the same texel (REPEAT)
    for (x, y) in [(0, 0), (6, 4), (3, 2)]:
        s, t = (x + 0.5) / W, (y + 0.5) / H
        got = oracle.tex_bilinear(img, [s, s + 1.0, s - 2.0], [t, t - 1.0, t + 3.0])
        want = (img[y, x, :3].astype(np.float32) / np.float32(255.0))
        assert np.array_equal(got[0], want)
        assert np.allclose(got[1:], want, atol=2e-6)
    # half way between the last and the first column: the mean of both (wrap-around)
    got = oracle.tex_bilinear(img, [0.0], [(2 + 0.5) / H])[0]
    assert np.allclose(got, (img[2, 6, :3] / 255.0 + img[2, 0, :3] / 255.0) / 2, atol=1e-6)
    # random coordinates against float64
    s = rng.uniform(-3, 3, 500).astype(np.float32)
    t = rng.uniform(-3, 3, 500).astype(np.float32)
    got = oracle.tex_bilinear(img, s, t)
    want = np.array([numpy_bilinear(img, a, b) for a, b in zip(s, t)])
    assert np.abs(got - want).max() < 5e-6
    # NaN / inf coordinates are black, never a crash
    assert (oracle.tex_bilinear(img, [np.nan, np.inf, 0.5], [0.5, 0.5, -np.inf]) == 0).all()


def test_barycentric_texture_coordinate(oracle):
    rng = np.random.default_rng(8)
    for _ in range(50):
        v = rng.uniform(-5, 5, (3, 3))
        uv = rng.uniform(0, 1, (3, 2))
        w = rng.dirichlet([1, 1, 1])
        p = w @ v
        got = oracle.tex_uv(v[0], v[1], v[2], p, uv.reshape(6))
        assert np.allclose(got, w @ uv, atol=2e-5)
    # at a corner: that corner's coordinate
    v = np.array([[0, 0, 0], [2, 0, 0], [0, 3, 0]], np.float32)
    uv = np.array([[0.1, 0.2], [0.9, 0.3], [0.4, 0.8]], np.float32)
    for k in range(3):
        assert np.allclose(oracle.tex_uv(v[0], v[1], v[2], v[k], uv.reshape(6)), uv[k], atol=1e-6)


def textured_scene(tmp_path):
    """The airplane stand-in with texture coordinates from the vertex positions and two of its six materials textured."""
    name, _ = scenes.write_airplane_standin(str(tmp_path), detail=24)
    scene = scenes.load_obj_scene(str(tmp_path), name, str(tmp_path))
    assert (scene["vertices"]["texture"] == 0).all()  # Q6: the loader never honours vt by default
    P = scene["vertices"]["vertex"]
    scene["vertices"]["texture"][:, 0] = P[:, 0] * np.float32(0.37) + P[:, 2] * np.float32(0.11)
    scene["vertices"]["texture"][:, 1] = P[:, 1] * np.float32(0.29) - P[:, 2] * np.float32(0.05)
    scene["materials"]["use_texture"][[0, 3]] = 1
    rng = np.random.default_rng(21)
    images = [(0, rng.integers(0, 256, (16, 32, 4), dtype=np.uint8)), (3, rng.integers(0, 256, (9, 5, 4), dtype=np.uint8))]
    return scene, images


@pytest.mark.gpu
@pytest.mark.parametrize("schedule", [0, 3])
def test_textured_render_matches_oracle_bit_exact(pkg, ctx, oracle, tmp_path, schedule):
    from test_parity_gpu import assert_bit_exact, cpu_render, gpu_render
    scene, images = textured_scene(tmp_path)
    W, H = 256, 128
    # option off: textures are ignored, the picture is the Kd picture (decree Q7)
    oracle.set_textures([])
    plain, _, oids, _ = cpu_render(oracle, W, H, True, 2, scene=scene)
    acc, _, ids, _ = gpu_render(pkg, ctx, W, H, True, 2, pkg.TRAV_REFERENCE, scene=scene, schedule=schedule, textures=images)
    assert_bit_exact(acc, plain)
    # option on
    try:
        oracle.set_textures(images)
        want, wout, _, wcnt = cpu_render(oracle, W, H, True, 2, scene=scene)
    finally:
        oracle.set_textures([])
    ctx.set_option(pkg.OPT_TEXTURES, 1)
    got, gout, gids, gcnt = gpu_render(pkg, ctx, W, H, True, 2, pkg.TRAV_REFERENCE, scene=scene, schedule=schedule, textures=images)
    assert np.array_equal(gids, oids)
    assert_bit_exact(got, want)
    assert np.array_equal(gout, wout)
    assert gcnt["segments"] == wcnt["segments"]
    assert not np.array_equal(want, plain)  # the textures are in the picture


@pytest.mark.gpu
def test_textures_need_schedule_0_or_3(pkg, ctx, tmp_path):
    scene, images = textured_scene(tmp_path)
    ctx.set_option(pkg.OPT_TEXTURES, 1)
    ctx.set_option(pkg.OPT_SCHEDULE, 2)
    ctx.upload_scene(scene)
    ctx.set_texture(images[0][0], images[0][1])
    unit, uniform = scenes.noise_tables(64, 32)
    ctx.upload_noise(unit, uniform)
    ctx.set_lights(scenes.model_lights())
    ctx.resize(64, 32)
    ctx.set_params(pkg.make_params(64, 32, scenes.reset_camera(True), True, accum_frames=1, bvh_count=1))
    with pytest.raises(RuntimeError, match="SRT_OPT_TEXTURES"):
        ctx.render(1)


@pytest.mark.gpu
def test_setting_a_materials_texture_again_replaces_it(pkg, ctx, oracle, tmp_path):
    """srt_set_texture for a material that already has an image: the new image replaces the old one (of another size, and
    with an image of a LATER material behind it in the texel buffer), nothing is appended beside it."""
    from test_parity_gpu import assert_bit_exact, cpu_render, gpu_render
    scene, images = textured_scene(tmp_path)
    W, H = 192, 96
    rng = np.random.default_rng(5)
    first = [(0, rng.integers(0, 256, (7, 3, 4), dtype=np.uint8)), images[1]]  # material 0's first image, then material 3's
    ctx.set_option(pkg.OPT_TEXTURES, 1)
    gpu_render(pkg, ctx, W, H, True, 1, pkg.TRAV_REFERENCE, scene=scene, schedule=3, textures=first)
    ctx.set_texture(images[0][0], images[0][1])  # replace material 0's image by a larger one
    ctx.resize(W, H)
    ctx.set_params(pkg.make_params(W, H, scenes.reset_camera(True), True, accum_frames=1, bvh_count=1))
    ctx.render(2)
    got = ctx.read_accum()
    try:
        oracle.set_textures(images)
        want, _, _, _ = cpu_render(oracle, W, H, True, 2, scene=scene)
    finally:
        oracle.set_textures([])
    assert_bit_exact(got, want)
