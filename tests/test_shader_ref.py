"""The hand-written oracle against the reference's own shader text, bit for bit (CPU).

oracle/_ref/libshader_ref.so is /root/reference/shaders/{raytrace_compute,raytrace_types,ray_intersects,raytrace_utils,
brdf}.glsl compiled as C++ where they lie (oracle/glsl_to_cpp.py: 14 syntactic rewrite rules; oracle/shim/glsl/glsl_shim.h:
vec types and built-ins forwarding to include/srt_detmath.h; oracle/shader_ref.cpp: the class the text is compiled into).
oracle/shader_oracle.cpp is the restatement every GPU parity test compares with.  Equal bits here -- accumulation image,
RGBA8 image, primary ids, segment / shadow-ray / box-test / triangle-test / accepted-hit / noise-fetch counts -- mean the
restatement reads the GLSL the way a compiler does, on the BASELINE.json configs the oracle can finish in seconds:
configs[0] exactly (spheres 640x480x1), Rubik (configs[1]'s scene) at 640x360x4, and the 20 k-triangle soup at depth 8
(configs[4]'s construction).  What both sides still share ON PURPOSE is the decree for what GLSL leaves undefined
(srt_detmath.h).
"""
import os
import tempfile

import numpy as np
import pytest

import scenes

REF_PRESENT = os.path.isdir("/root/reference/shaders")


@pytest.fixture(scope="module")
def both(oracle):
    if not oracle.shader_ref_available():
        if REF_PRESENT:
            pytest.fail("oracle/_ref/libshader_ref.so is missing although /root/reference is here: run `make -C oracle all`")
        pytest.skip("oracle/_ref/libshader_ref.so not built (no /root/reference on this machine)")
    return oracle


def bits(a):
    a = np.ascontiguousarray(a, np.float32)
    return np.where(np.isnan(a), np.uint32(0x7FC00000), a.view(np.uint32))


def compare(O, W, H, first, frames, show_model, scene, cam, lights, max_depth=0, bvh_count=1, noise_seed=0x5EED):
    unit, uniform = scenes.noise_tables(W, H, seed=noise_seed)
    p = O.make_params(W, H, cam, show_model, first, len(lights), bvh_count if show_model else 0, max_depth)
    a = O.render(p, frames, scene, unit, uniform, lights)
    b = O.render(p, frames, scene, unit, uniform, lights, impl="shader_ref")
    assert np.array_equal(a[2], b[2]), "primary ids differ"
    assert np.array_equal(bits(a[0]), bits(b[0])), "accumulation images differ"
    assert np.array_equal(a[1], b[1]), "RGBA8 images differ"
    for k in ("segments", "shadow_rays", "nodes", "tris", "accepted", "noise_lookups", "capped_paths", "path_len_hist"):
        assert a[3][k] == b[3][k], k
    assert a[3]["capped_paths"] == 0
    return a[3]


def test_config0_spheres_640x480_1spp(both):
    cnt = compare(both, 640, 480, 1, 1, False, None, scenes.reset_camera(False), scenes.sphere_lights())
    assert cnt["segments"] > 640 * 480 and cnt["shadow_rays"] > 0


def test_spheres_later_frames_and_other_noise(both):
    compare(both, 320, 240, 7, 3, False, None, scenes.reset_camera(False), scenes.sphere_lights(), noise_seed=77)


def test_rubik_640x360_4spp(both):
    cnt = compare(both, 640, 360, 1, 4, True, scenes.rubik_scene(), scenes.reset_camera(True), scenes.model_lights())
    assert cnt["tris"] > cnt["nodes"] > cnt["segments"] > 0


def test_rubik_from_inside_depth8(both):
    cam = scenes.look_camera((0.3, 9.1, 12.0), -70.0, -12.0)
    compare(both, 160, 120, 1, 2, True, scenes.rubik_scene(), cam, scenes.model_lights(), max_depth=8)


def test_soup_20k_depth8(both):
    with tempfile.TemporaryDirectory() as d:
        soup, info = scenes.soup_scene(20000, d)
    cam = scenes.look_camera((0.0, 0.0, 120.0), -90.0, 0.0)
    cnt = compare(both, 320, 240, 1, 2, True, soup, cam, scenes.model_lights(), max_depth=8)
    assert cnt["path_len_hist"][1] < 320 * 240 * 2  # some camera rays do hit the cloud


def test_two_models_with_frames(both):
    """bvh_count = 2 (raytrace_compute.glsl:144-161): the second model behind a translated frame."""
    a = scenes.rubik_scene()
    b = scenes.rubik_scene()
    two = scenes.concat_scenes(a, b)
    two["bvhs"]["frame"][:] = np.eye(4, dtype=np.float32).reshape(-1)
    two["bvhs"]["frame"][1][12:15] = (-14.0, 3.0, 6.0)  # world -> model translation, column-major
    compare(both, 256, 144, 1, 2, True, two, scenes.reset_camera(True), scenes.model_lights(), bvh_count=2)


def test_reset_dispatch(both):
    W, H = 64, 48
    unit, uniform = scenes.noise_tables(W, H)
    lights, cam = scenes.sphere_lights(), scenes.reset_camera(False)
    p = both.make_params(W, H, cam, False, 1, len(lights), 0)
    acc, _, _, _ = both.render(p, 2, None, unit, uniform, lights, impl="shader_ref")
    pr = both.make_params(W, H, cam, False, 1, len(lights), 0, reset=1)
    acc, _, _, cnt = both.render(pr, 1, None, unit, uniform, lights, accum=acc, impl="shader_ref")
    assert not acc[..., :3].any() and (acc[..., 3] == 1.0).all() and cnt["segments"] == 0
