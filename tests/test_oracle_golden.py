"""CPU tests of the oracle itself: committed golden frames, schedule independence, continuation.

The oracle (oracle/shader_oracle.cpp) is the checker of every GPU parity test, so it is pinned here first: against
tests/golden/shader_ref_frames.npz -- frames rendered by THE REFERENCE'S OWN SHADER TEXT compiled as C++
(oracle/_ref/libshader_ref.so, made from /root/reference/shaders/*.glsl by oracle/Makefile; fixtures written by
tests/golden/make_shader_ref_golden.py so that this test also runs where /root/reference is absent) -- and against
properties the shader guarantees: pixels are independent (any row split / thread count gives the same bits) and
frames accumulate in order (rendering frames 1..2 then 3..4 on the same buffer equals rendering 1..4).
tests/test_shader_ref.py makes the same comparison live, at larger sizes, where the compiled shader is present.
"""
import os
import sys

import numpy as np
import pytest

import scenes

sys.path.insert(0, os.path.join(scenes.ROOT, "tests", "golden"))
import make_shader_ref_golden as G  # noqa: E402


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(scenes.GOLDEN, "shader_ref_frames.npz"))


def bits(a):
    a = np.ascontiguousarray(a, np.float32)
    return np.where(np.isnan(a), np.uint32(0x7FC00000), a.view(np.uint32))


@pytest.mark.parametrize("name", sorted(G.CASES))
def test_oracle_reproduces_golden_frames(oracle, golden, name):
    acc, out, ids, cnt = G.render_case(name)
    assert np.array_equal(ids, golden[name + "/ids"])
    assert np.array_equal(bits(acc), bits(golden[name + "/accum"]))
    assert np.array_equal(out, golden[name + "/out"])
    assert [cnt[k] for k in G.COUNTERS] == golden[name + "/rays"].tolist()
    assert cnt["capped_paths"] == 0


def _rubik_inputs(W, H):
    unit, uniform = scenes.noise_tables(W, H)
    return scenes.rubik_scene(), unit, uniform, scenes.model_lights(), scenes.reset_camera(True)


def test_threads_and_row_bands_do_not_change_bits(oracle):
    W, H = 96, 56
    scene, unit, uniform, lights, cam = _rubik_inputs(W, H)
    p = oracle.make_params(W, H, cam, True, 1, len(lights), 1)
    whole, out, ids, cnt = oracle.render(p, 2, scene, unit, uniform, lights, n_threads=1)
    threaded, out_t, ids_t, cnt_t = oracle.render(p, 2, scene, unit, uniform, lights, n_threads=5)
    assert np.array_equal(bits(whole), bits(threaded)) and np.array_equal(ids, ids_t) and np.array_equal(out, out_t)
    assert cnt["segments"] == cnt_t["segments"] and cnt["tris"] == cnt_t["tris"]
    banded = np.zeros_like(whole)
    for y0, y1 in ((0, 16), (16, 24), (24, 56)):
        oracle.render(p, 2, scene, unit, uniform, lights, accum=banded, rows=(y0, y1), n_threads=2)
    assert np.array_equal(bits(whole), bits(banded))


def test_frames_continue_on_the_same_accumulation_buffer(oracle):
    W, H = 64, 48
    unit, uniform = scenes.noise_tables(W, H)
    lights, cam = scenes.sphere_lights(), scenes.reset_camera(False)
    p = oracle.make_params(W, H, cam, False, 1, len(lights), 0)
    four, out4, _, _ = oracle.render(p, 4, None, unit, uniform, lights)
    acc, _, _, _ = oracle.render(p, 2, None, unit, uniform, lights)
    p2 = oracle.make_params(W, H, cam, False, 3, len(lights), 0)
    acc, out2, _, _ = oracle.render(p2, 2, None, unit, uniform, lights, accum=acc)
    assert np.array_equal(bits(four), bits(acc)) and np.array_equal(out4, out2)
    assert (four[..., 3] == 1.0).all()  # accum.w is always 1 (raytrace_compute.glsl:406)


def test_reset_frame_clears_and_renders_nothing(oracle):
    W, H = 32, 24
    unit, uniform = scenes.noise_tables(W, H)
    lights, cam = scenes.sphere_lights(), scenes.reset_camera(False)
    p = oracle.make_params(W, H, cam, False, 1, len(lights), 0)
    acc, _, _, _ = oracle.render(p, 1, None, unit, uniform, lights)
    assert acc[..., :3].any()
    pr = oracle.make_params(W, H, cam, False, 1, len(lights), 0, reset=1)
    acc, _, _, cnt = oracle.render(pr, 1, None, unit, uniform, lights, accum=acc)
    assert not acc[..., :3].any() and (acc[..., 3] == 1.0).all() and cnt["segments"] == 0
