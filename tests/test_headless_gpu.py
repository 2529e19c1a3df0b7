"""GPU test of the host front end: srt_headless -- the reference's frame loop on the
reference-shaped C++ classes (Compute, Texture, Camera, LoadObject, UploadModelDataToGPU) -- must put the same bits
into the accumulation image as the oracle does for the same OBJ, noise tables, lights and camera; and building the
model's BVH on the device (--device-bvh) must not change a byte.
"""
import json
import os
import subprocess

import numpy as np
import pytest

import scenes
from test_parity_gpu import assert_bit_exact, cpu_render

pytestmark = pytest.mark.gpu

HOST = os.path.join(scenes.ROOT, "simple-ray-tracer_b200", "host")


def run_headless(tmp_path, name, W, H, frames, tag, extra=()):
    exe = os.path.join(HOST, "srt_headless")
    if not os.path.exists(exe):
        subprocess.run(["make", "-s", "-C", HOST, "srt_headless"], check=True)
    accum, ppm = str(tmp_path / (tag + ".accum")), str(tmp_path / (tag + ".ppm"))
    r = subprocess.run([exe, "--model", name, "--objects", str(tmp_path) + "/", "--width", str(W), "--height", str(H),
                        "--frames", str(frames), "--batch", "2", "--noise", str(tmp_path / "noise.bin"), "--accum", accum,
                        "--out", ppm, *extra], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    info = json.loads(r.stdout.strip().splitlines()[-1])
    return np.fromfile(accum, np.float32).reshape(H, W, 4), open(ppm).read(), info


def test_headless_front_end_matches_oracle_and_device_bvh_changes_nothing(oracle, tmp_path):
    W, H, frames = 256, 128, 3
    name, n_tri = scenes.write_airplane_standin(str(tmp_path), detail=32)
    unit, uniform = scenes.noise_tables(W, H)
    with open(tmp_path / "noise.bin", "wb") as f:
        f.write(unit.tobytes())
        f.write(uniform.tobytes())
    acc_host, ppm_host, info_host = run_headless(tmp_path, name, W, H, frames, "host")
    acc_dev, ppm_dev, info_dev = run_headless(tmp_path, name, W, H, frames, "dev", ["--device-bvh"])
    assert info_host["bvh"] == "host" and info_dev["bvh"] == "device"
    assert acc_host.tobytes() == acc_dev.tobytes() and ppm_host == ppm_dev
    assert info_host["segments"] == info_dev["segments"]

    scene = scenes.load_obj_scene(str(tmp_path), name, str(tmp_path))
    assert len(scene["triangles"]) == n_tri
    oacc, oout, _, ocnt = cpu_render(oracle, W, H, True, frames, scene=scene, noise=(unit, uniform))
    assert_bit_exact(acc_host, oacc)
    assert info_host["segments"] == ocnt["segments"] and info_host["shadow_rays"] == ocnt["shadow_rays"]
    # the PPM (P3) holds the RGBA8 picture the oracle tonemaps, top row first (raytracer.cpp WritePPM)
    tok = ppm_host.split()
    assert tok[0] == "P3" and int(tok[1]) == W and int(tok[2]) == H and int(tok[3]) == 255
    rgb = np.array(tok[4:], np.int64).reshape(H, W, 3)
    want = oout[..., :3].astype(np.int64)
    assert np.array_equal(rgb, want[::-1])


def write_textured_obj(tmp_path, name="wall"):
    """A textured wall (two quads with vt, map_Kd) over an untextured floor, in front of the model camera."""
    d = tmp_path / name
    d.mkdir()
    rng = np.random.default_rng(17)
    img = rng.integers(0, 256, (8, 16, 3), dtype=np.uint8)
    scenes.write_png(d / "wall_diff.png", img, filters=[0, 1, 2, 3, 4])
    (d / (name + ".mtl")).write_text(
        "newmtl painted\nKd 0.3 0.3 0.3\nKs 0.2 0.2 0.2\nNs 20\nmap_Kd wall_diff.png\n"
        "newmtl plain\nKd 0.7 0.6 0.5\nKs 0.1 0.1 0.1\nNs 5\n")
    v = [(-12, 0, 0), (0, 0, -3), (12, 0, 0), (12, 18, 0), (0, 18, -3), (-12, 18, 0),  # wall, folded
         (-20, 0, 20), (20, 0, 20), (20, 0, -5), (-20, 0, -5)]                          # floor
    vt = [(0.0, 0.0), (1.3, 0.0), (2.6, 0.0), (2.6, 1.7), (1.3, 1.7), (0.0, 1.7)]
    lines = ["mtllib %s.mtl" % name] + ["v %g %g %g" % p for p in v] + ["vt %g %g" % t for t in vt]
    lines += ["usemtl painted", "f 1/1 2/2 5/5 6/6", "f 2/2 3/3 4/4 5/5", "usemtl plain", "f 7 8 9 10"]
    (d / (name + ".obj")).write_text("\n".join(lines) + "\n")
    return name, np.concatenate([img, np.full((8, 16, 1), 255, np.uint8)], axis=2)


def test_headless_textures_option_matches_oracle(oracle, tmp_path):
    """SURVEY.md 8(f) N3 end to end: OBJ with vt + MTL with map_Kd -> host loader (PNG decode, texture coordinates) ->
    srt_set_texture / SRT_OPT_TEXTURES -> picture, against the oracle fed the same arrays and the pixels the PNG was
    written from.  Without --textures the same files give the Kd picture (decree Q7)."""
    W, H, frames = 256, 128, 2
    name, pixels = write_textured_obj(tmp_path)
    unit, uniform = scenes.noise_tables(W, H)
    with open(tmp_path / "noise.bin", "wb") as f:
        f.write(unit.tobytes())
        f.write(uniform.tobytes())
    acc_plain, _, _ = run_headless(tmp_path, name, W, H, frames, "plain")
    acc_tex, _, _ = run_headless(tmp_path, name, W, H, frames, "tex", ["--textures"])
    assert not np.array_equal(acc_plain, acc_tex)

    plain_scene = scenes.load_obj_scene(str(tmp_path), name, str(tmp_path))
    assert (plain_scene["vertices"]["texture"] == 0).all()  # Q6
    oracle.set_textures([])
    oacc, _, _, _ = cpu_render(oracle, W, H, True, frames, scene=plain_scene, noise=(unit, uniform))
    assert_bit_exact(acc_plain, oacc)

    scene = scenes.load_obj_scene(str(tmp_path), name, str(tmp_path), textures=True)
    assert (scene["vertices"]["texture"] != 0).any()
    painted = int(np.nonzero(scene["materials"]["use_texture"])[0][0])
    try:
        oracle.set_textures([(painted, pixels)])
        oacc_tex, _, _, _ = cpu_render(oracle, W, H, True, frames, scene=scene, noise=(unit, uniform))
    finally:
        oracle.set_textures([])
    assert_bit_exact(acc_tex, oacc_tex)
