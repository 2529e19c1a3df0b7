"""Generates tests/golden/host_inputs.json from THE REFERENCE'S OWN host code (oracle/_ref/ref_host_inputs, compiled by
oracle/Makefile from /root/reference/include/common/utils.h and src/raytracer/camera.cpp where they lie):
  * the first NOISE_N texels of both noise tables as UpdateNoiseTex fills them (src/main.cpp:269-279, unseeded std::rand()
    = glibc's srand(1) stream) -- SURVEY.md 8a F24;
  * the camera's origin/front/up/right after every line of a fixed script of Reset / Rotate / Move* / MoveAndRotate calls
    (src/raytracer/camera.cpp) -- F25.
tests/test_host_inputs.py holds simple-ray-tracer_b200/host to these bytes on any machine.  Run from the repo root:

    make -C oracle all && python tests/golden/make_host_inputs_golden.py
"""
import json
import os
import random
import subprocess
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF_TOOL = os.path.join(ROOT, "oracle", "_ref", "ref_host_inputs")
NOISE_N = 512


def camera_script(seed=3, n=400):
    rnd = random.Random(seed)
    lines = ["reset 1"]
    for _ in range(n):
        r = rnd.random()
        if r < 0.35:
            lines.append("rotate %.4f %.4f" % (rnd.uniform(-40, 40), rnd.uniform(-30, 30)))
        elif r < 0.7:
            lines.append("%s %.4f" % (rnd.choice("forward backward left right up down".split()), rnd.uniform(0, 3)))
        elif r < 0.95:
            lines.append("moverot %.4f %.3f %.3f %.3f %.3f %.3f %.3f" % (
                rnd.uniform(0.001, 0.05), rnd.uniform(-1, 1), rnd.uniform(-1, 1), rnd.uniform(-1, 1), rnd.uniform(-5, 5),
                rnd.uniform(-5, 5), rnd.uniform(1, 20)))
        else:
            lines.append("reset %d" % rnd.randint(0, 1))
    return lines


def run_tool(tool, script_lines, noise_n):
    noise = subprocess.run([tool, "noise", str(noise_n)], check=True, capture_output=True, text=True).stdout.split()
    with tempfile.NamedTemporaryFile("w", suffix=".txt", delete=False) as f:
        f.write("\n".join(script_lines) + "\n")
    try:
        cam = subprocess.run([tool, "camera", f.name], check=True, capture_output=True, text=True).stdout.splitlines()
    finally:
        os.remove(f.name)
    return noise, [ln.split() for ln in cam]


if __name__ == "__main__":
    script = camera_script()
    noise, cam = run_tool(REF_TOOL, script, NOISE_N)
    assert len(noise) == 2 * NOISE_N * 3 and len(cam) == len(script)
    path = os.path.join(ROOT, "tests", "golden", "host_inputs.json")
    with open(path, "w") as f:
        json.dump({"noise_n": NOISE_N, "noise_hex": noise, "camera_script": script, "camera_hex": cam}, f)
    print(path, os.path.getsize(path), "bytes; first unit vector", noise[:3])
