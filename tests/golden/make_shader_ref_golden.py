"""Generates tests/golden/shader_ref_frames.npz: small frames rendered by THE REFERENCE'S OWN SHADER TEXT.

oracle/_ref/libshader_ref.so is /root/reference/shaders/*.glsl compiled as C++ where the files lie (oracle/Makefile,
oracle/glsl_to_cpp.py, oracle/shader_ref.cpp over oracle/shim/glsl/glsl_shim.h; no GL, no GLSL compiler).  It exists
only where /root/reference does, so its outputs for the cases below are committed as fixtures: tests/test_oracle_golden.py
holds the hand-written oracle (oracle/shader_oracle.cpp, the checker of every GPU parity test) to these bits on any
machine, and the GPU tests compare the CUDA path with the same files.  Inputs are the seeded tables of tests/scenes.py.
Run from the repo root, in the container that has /root/reference:

    make -C oracle all && python tests/golden/make_shader_ref_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import scenes  # noqa: E402
from oracle import oracle_py as O  # noqa: E402

CASES = {
    # name: (show_model, W, H, first_frame, n_frames, max_depth, camera)
    "spheres_64x48_f1": (False, 64, 48, 1, 1, 0, None),
    "spheres_64x48_f3_2": (False, 64, 48, 3, 2, 0, None),
    "spheres_160x120_f1_3": (False, 160, 120, 1, 3, 0, None),
    "rubik_96x56_f1_2": (True, 96, 56, 1, 2, 0, None),
    "rubik_192x104_f1_4": (True, 192, 104, 1, 4, 0, None),
    "rubik_inside_48x48_d8": (True, 48, 48, 1, 2, 8, ((0.3, 9.1, 12.0), -70.0, -12.0)),
    "rubik_close_64x64_f5_2": (True, 64, 64, 5, 2, 0, ((6.0, 14.0, 22.0), -105.0, -15.0)),
}
COUNTERS = ("segments", "shadow_rays", "nodes", "tris", "accepted", "noise_lookups")


def render_case(name, impl="oracle"):
    show_model, W, H, first, n, depth, pose = CASES[name]
    unit, uniform = scenes.noise_tables(W, H)
    lights = scenes.model_lights() if show_model else scenes.sphere_lights()
    cam = scenes.look_camera(*pose) if pose else scenes.reset_camera(show_model)
    scene = scenes.rubik_scene() if show_model else None
    p = O.make_params(W, H, cam, show_model, first, len(lights), 1 if show_model else 0, depth)
    acc, out, ids, cnt = O.render(p, n, scene, unit, uniform, lights, n_threads=1, impl=impl)
    return acc, out, ids, cnt


if __name__ == "__main__":
    blob = {}
    for name in CASES:
        acc, out, ids, cnt = render_case(name, impl="shader_ref")
        blob[name + "/accum"] = acc
        blob[name + "/out"] = out
        blob[name + "/ids"] = ids
        blob[name + "/rays"] = np.array([cnt[k] for k in COUNTERS], np.uint64)
        print(name, [cnt[k] for k in COUNTERS], "capped", cnt["capped_paths"])
    path = os.path.join(ROOT, "tests", "golden", "shader_ref_frames.npz")
    np.savez_compressed(path, **blob)
    print(path, os.path.getsize(path), "bytes")
