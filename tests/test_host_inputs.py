"""The host-made inputs of the hot path (SURVEY.md 8a F24 noise tables, F25 camera basis) against the reference's own code.

The product's host classes (simple-ray-tracer_b200/host: Common::randomVec3 / randomUnitVector / fillNoiseTables,
RayTracer::Camera) must hand the kernels the very floats the reference's host code hands its shader.  Checked twice:
against tests/golden/host_inputs.json (written by tests/golden/make_host_inputs_golden.py from the reference's
include/common/utils.h and src/raytracer/camera.cpp compiled in oracle/_ref) on any machine, and live against
oracle/_ref/ref_host_inputs with a longer table where that binary exists.
"""
import json
import os
import sys

import pytest

import scenes

sys.path.insert(0, os.path.join(scenes.ROOT, "tests", "golden"))
import make_host_inputs_golden as G  # noqa: E402


@pytest.fixture(scope="module")
def probe():
    return scenes._host_tool("srt_host_probe")


@pytest.fixture(scope="module")
def golden():
    with open(os.path.join(scenes.GOLDEN, "host_inputs.json")) as f:
        return json.load(f)


def test_noise_tables_equal_the_reference_fill(probe, golden):
    noise, _ = G.run_tool(probe, ["reset 1"], golden["noise_n"])
    assert noise == golden["noise_hex"]
    # after srand(1) glibc's first three draws are 0.8402 / 0.3944 / 0.7831; g++ evaluates the reference's
    # glm::vec3(randomFloat(), randomFloat(), randomFloat()) right to left, so the first draw lands in z
    import struct
    n = golden["noise_n"]
    first_uniform = [struct.unpack("<f", bytes.fromhex(h)[::-1])[0] for h in golden["noise_hex"][3 * n:3 * n + 3]]
    assert all(0.0 <= v < 1.0 for v in first_uniform)


def test_camera_basis_equals_the_reference_camera(probe, golden):
    _, cam = G.run_tool(probe, golden["camera_script"], 1)
    assert len(cam) == len(golden["camera_hex"])
    for i, (a, b) in enumerate(zip(cam, golden["camera_hex"])):
        assert a == b, "line %d (%s)" % (i, golden["camera_script"][i])
    # Reset with the model scene: origin (0,9,40), front.x = (float)cos(-pi/2 as a double) = -4.371139e-08
    assert golden["camera_hex"][0][:6] == ["00000000", "41100000", "42200000", "b33bbd2e", "00000000", "bf800000"]


def test_live_against_the_reference_binary(probe):
    if not os.path.exists(G.REF_TOOL):
        if os.path.isdir("/root/reference/src"):
            pytest.fail("oracle/_ref/ref_host_inputs missing although /root/reference is here: run `make -C oracle all`")
        pytest.skip("oracle/_ref/ref_host_inputs not built (no /root/reference on this machine)")
    script = G.camera_script(seed=11, n=600)
    ref_noise, ref_cam = G.run_tool(G.REF_TOOL, script, 20000)
    noise, cam = G.run_tool(probe, script, 20000)
    assert noise == ref_noise
    assert cam == ref_cam
