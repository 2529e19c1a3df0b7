"""Host-only check of the hit finder (simple-ray-tracer_b200/csrc/srt_finder.cu) and of the algorithm its kernels run:
tests/native/finder_emulation.cu builds the finder for a scene, emulates the candidate walk + the replay of the shader's
box culls (and the distance-bounded walk for rays that start outside the root box) and compares every ray with an
emulation of the shader's own traversal (ray_intersects.glsl:99-133).  No GPU: the CUDA kernels are compared with the
oracle bit for bit in tests/test_parity_gpu.py."""
import os
import re
import shutil
import subprocess

import numpy as np
import pytest

import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def emulator(tmp_path_factory):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not found (the finder's records are the device's: srt_finder.cu is compiled by nvcc)")
    out = str(tmp_path_factory.mktemp("finder") / "finder_emulation")
    csrc = os.path.join(ROOT, "simple-ray-tracer_b200", "csrc")
    subprocess.run([nvcc, "-O2", "-std=c++17", "-Wno-deprecated-gpu-targets", "-I" + os.path.join(ROOT, "include"), "-I" + csrc,
                    "-o", out, os.path.join(ROOT, "tests", "native", "finder_emulation.cu"), os.path.join(csrc, "srt_finder.cu"),
                    # build_finder can hand its trees to the device builder (not here: no workspace is passed), which links
                    # against these two
                    os.path.join(csrc, "srt_finder_build.cu"), os.path.join(csrc, "srt_bvh_build.cu")],
                   check=True)
    return out


def run(emulator, scene, tmp):
    for k, f in (("bvhs", "bvhs"), ("nodes", "nodes"), ("triangles", "tris"), ("vertices", "verts")):
        np.ascontiguousarray(scene[k]).tofile(os.path.join(tmp, f + ".bin"))
    p = subprocess.run([emulator, tmp], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout + p.stderr
    m = re.search(r"outside rays (\d+): mismatches (\d+) uncertain (\d+) overflow (\d+)", p.stdout)
    n = re.search(r"rays (\d+) mismatches (\d+) overflow", p.stdout)
    assert m and n, p.stdout
    return [int(v) for v in m.groups()], [int(v) for v in n.groups()], p.stdout


@pytest.mark.parametrize("which", ["rubik", "airplane", "soup"])
def test_finder_and_outside_walk_agree_with_the_shaders_traversal(emulator, tmp_path, which):
    tmp = str(tmp_path)
    if which == "rubik":
        scene = scenes.rubik_scene()
    elif which == "airplane":
        name, _ = scenes.write_airplane_standin(tmp, detail=24)
        scene = scenes.load_obj_scene(tmp, name, tmp)
    else:
        scene, _ = scenes.soup_scene(5000, tmp)
    (n_out, out_mism, out_unsure, out_over), (n_rays, mism), text = run(emulator, scene, tmp)
    assert n_rays == 200000 and mism == 0, text          # candidates + replay == the shader's walk, hit and distance
    assert n_out > 50000 and out_mism == 0, text         # the distance-bounded walk too, on every ray it answers itself
    assert out_unsure + out_over < 0.01 * n_out, text    # and it hands (almost) none to the long way
