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


def _camera_after(script_lines):
    """Camera vectors after every line of a srt_host_probe script (the product's RayTracer::Camera, which
    tests/test_host_inputs.py holds to the reference's camera.cpp bit for bit)."""
    probe = scenes._host_tool("srt_host_probe")
    path = os.path.join(scenes.ROOT, "tests", ".probe_script.tmp")
    with open(path, "w") as f:
        f.write("\n".join(script_lines) + "\n")
    try:
        out = subprocess.run([probe, "camera", path], check=True, capture_output=True, text=True).stdout.splitlines()
    finally:
        os.remove(path)
    cams = []
    for ln in out:
        v = np.array([int(h, 16) for h in ln.split()], np.uint32).view(np.float32)
        cams.append({"origin": tuple(map(float, v[0:3])), "direction": tuple(map(float, v[3:6])), "up": tuple(map(float, v[6:9])),
                     "right": tuple(map(float, v[9:12]))})
    return cams


def test_scripted_camera_path_with_the_references_reset_semantics(oracle, tmp_path):
    """--script replaces the InputHandler.  Any input raises resetAccumBuffer and sets
    accumFrames = 0 before the frame's accumFrames++ (src/main.cpp:629-659), so that dispatch only clears with
    accumFrames = 1 (raytrace_compute.glsl:390-393) and the first sample after it is taken at accumFrames = 2 and divided
    by 2.  The oracle is driven through the same sequence of (reset, accumFrames, camera) dispatches."""
    W, H = 192, 104
    name, _ = scenes.write_airplane_standin(str(tmp_path), detail=24)
    unit, uniform = scenes.noise_tables(W, H)
    with open(tmp_path / "noise.bin", "wb") as f:
        f.write(unit.tobytes())
        f.write(uniform.tobytes())
    mid = str(tmp_path / "mid.ppm")
    script = tmp_path / "path.txt"
    script.write_text("frames 3\n"
                      "move 0.016 0.5 0 1 4.0 -2.0   # walk forward/right while turning: a reset frame\n"
                      "frames 2\n"
                      "write %s\n"
                      "click\n"
                      "frames 4\n" % mid)
    acc, ppm, info = run_headless(tmp_path, name, W, H, 0, "scripted", ["--script", str(script)])
    assert info["resets"] == 2 and info["accum_frames"] == 5 and info["frames"] == 3 + 1 + 2 + 1 + 4

    # the same camera, call by call: Reset, 3 idle MoveAndRotate calls, the move, 2 idle, the click (no deltas), 4 idle
    idle = "moverot 0.016 0 0 0 0 0 1"
    cams = _camera_after(["reset 1"] + [idle] * 3 + ["moverot 0.016 0.5 0 1 4.0 -2.0 1"] + [idle] * 2 + [idle] + [idle] * 4)
    cam0, cam1 = cams[0], cams[4]
    assert cams[3] == cam0 and cams[-1] == cam1 and cam1 != cam0
    scene = scenes.load_obj_scene(str(tmp_path), name, str(tmp_path))
    lights = scenes.model_lights()

    def run(accum, cam, first, n, reset=0):
        p = oracle.make_params(W, H, cam, True, first, len(lights), 1, 0, reset=reset)
        a, out, _, _ = oracle.render(p, n, scene, unit, uniform, lights, accum=accum)
        return a, out

    a, _ = run(None, cam0, 1, 3)                 # frames 3: accumFrames 1..3
    a, _ = run(a, cam1, 1, 1, reset=1)           # the move: clears, accumFrames = 1
    a, out_mid = run(a, cam1, 2, 2)              # frames 2: accumFrames 2, 3 (two samples divided by 3)
    a, _ = run(a, cam1, 1, 1, reset=1)           # the click: clears again
    a, out_end = run(a, cam1, 2, 4)              # frames 4: accumFrames 2..5 (four samples divided by 5)
    assert_bit_exact(acc, a)

    def ppm_rgb(text):
        tok = text.split()
        return np.array(tok[4:], np.int64).reshape(H, W, 3)

    assert np.array_equal(ppm_rgb(ppm), out_end[..., :3].astype(np.int64)[::-1])
    assert np.array_equal(ppm_rgb(open(mid).read()), out_mid[..., :3].astype(np.int64)[::-1])
    # Q11 made visible: four samples were taken since the click, and the picture divides their sum by five
    lin = np.clip(a[..., :3] / np.float32(4.0), 0, 1)
    brighter = np.where(lin < 0.0031308, lin * 12.92, 1.055 * np.power(lin, 1 / 2.4) - 0.055)
    shown = out_end[..., :3].astype(np.float64) / 255.0
    assert (shown <= brighter + 1.5 / 255).all() and (shown < brighter - 2.0 / 255).mean() > 0.5


def test_obj_option_is_the_objects_and_model_pair(oracle, tmp_path):
    W, H = 128, 64
    name, _ = scenes.write_airplane_standin(str(tmp_path), detail=16)
    unit, uniform = scenes.noise_tables(W, H)
    with open(tmp_path / "noise.bin", "wb") as f:
        f.write(unit.tobytes())
        f.write(uniform.tobytes())
    a, _, _ = run_headless(tmp_path, name, W, H, 2, "pair")
    exe = os.path.join(HOST, "srt_headless")
    accum = str(tmp_path / "obj.accum")
    r = subprocess.run([exe, "--obj", str(tmp_path / name / (name + ".obj")), "--width", str(W), "--height", str(H), "--frames", "2",
                        "--noise", str(tmp_path / "noise.bin"), "--accum", accum, "--out", str(tmp_path / "obj.ppm")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert np.fromfile(accum, np.float32).tobytes() == a.tobytes()
