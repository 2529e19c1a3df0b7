"""Generates tests/golden/oracle_frames.npz: small frames rendered by the CPU oracle (oracle/liboracle.so).

The reference's shader cannot run on this box and its own tests do not pin the path, so these
fixtures pin the ORACLE AGAINST ITSELF: any later edit of oracle/shader_oracle.cpp or include/srt_detmath.h that
changes a result shows up as a diff here (tests/test_oracle_golden.py), and the GPU tests compare the CUDA path with
the same files.  Inputs are the seeded tables of tests/scenes.py.  Run from the repo root:

    make -C oracle liboracle.so && python tests/golden/make_oracle_golden.py
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
    "rubik_96x56_f1_2": (True, 96, 56, 1, 2, 0, None),
    "rubik_inside_48x48_d8": (True, 48, 48, 1, 2, 8, ((0.3, 9.1, 12.0), -70.0, -12.0)),
}


def render_case(name):
    show_model, W, H, first, n, depth, pose = CASES[name]
    unit, uniform = scenes.noise_tables(W, H)
    lights = scenes.model_lights() if show_model else scenes.sphere_lights()
    cam = scenes.look_camera(*pose) if pose else scenes.reset_camera(show_model)
    scene = scenes.rubik_scene() if show_model else None
    p = O.make_params(W, H, cam, show_model, first, len(lights), 1 if show_model else 0, depth)
    acc, out, ids, cnt = O.render(p, n, scene, unit, uniform, lights, n_threads=1)
    return acc, out, ids, cnt


if __name__ == "__main__":
    blob = {}
    for name in CASES:
        acc, out, ids, cnt = render_case(name)
        blob[name + "/accum"] = acc
        blob[name + "/out"] = out
        blob[name + "/ids"] = ids
        blob[name + "/rays"] = np.array([cnt["segments"], cnt["shadow_rays"], cnt["nodes"], cnt["tris"]], np.uint64)
        print(name, cnt["segments"], cnt["shadow_rays"], cnt["nodes"], cnt["tris"], "nan", cnt["nan_pixels"])
    path = os.path.join(ROOT, "tests", "golden", "oracle_frames.npz")
    np.savez_compressed(path, **blob)
    print(path, os.path.getsize(path), "bytes")
