"""Deterministic inputs shared by the parity tests, smoke() and bench.py (no reference access at run time).

Everything here is an INPUT of both the oracle and the CUDA path: noise tables (src/main.cpp:269-279 fills
them with std::rand(); here a PCG64 stream so that they do not depend on the libc), the light lists of
src/main.cpp:582-595, the camera basis of RayTracer::Camera::Reset (src/raytracer/camera.cpp:187-212).
"""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

LIGHT_DT = np.dtype([("position", "<f4", 3), ("intensity", "<f4"), ("color", "<f4", 3), ("pad1", "<f4")])


def _lights(rows):
    a = np.zeros(len(rows), LIGHT_DT)
    for i, (pos, col, inten) in enumerate(rows):
        a[i]["position"], a[i]["color"], a[i]["intensity"] = pos, col, inten
    return a


def model_lights():  # src/main.cpp:584-589
    return _lights([((1.0, 10.0, 10.0), (1.0, 1.0, 1.0), 50.0), ((-5.0, 15.0, 10.0), (1.0, 0.2, 0.2), 15.0),
                    ((5.0, 15.0, 10.0), (0.2, 1.0, 0.2), 15.0), ((-5.0, 5.0, 10.0), (0.2, 0.2, 1.0), 15.0),
                    ((5.0, 5.0, 10.0), (1.0, 1.0, 0.1), 15.0), ((0.0, 21.0, 17.0), (1.0, 1.0, 1.0), 50.0)])


def sphere_lights():  # src/main.cpp:593-594
    return _lights([((1.0, 2.0, 0.0), (1.0, 1.0, 1.0), 10.0), ((-2.5, 2.0, 0.0), (1.0, 1.0, 1.0), 3.0)])


def reset_camera(show_model, origin=None):
    """Camera::Reset + UpdateCameraVectors with yaw=-90, pitch=0: cosf(radians(-90)) = -4.371139e-08."""
    c = np.float32(-4.371139e-08)
    if origin is None:
        origin = (0.0, 9.0, 40.0) if show_model else (0.0, 1.0, 4.0)
    return {"origin": tuple(float(v) for v in origin), "direction": (float(c), 0.0, -1.0), "up": (0.0, 1.0, 0.0),
            "right": (1.0, 0.0, float(c))}


def look_camera(origin, yaw_deg, pitch_deg):
    """Camera::UpdateCameraVectors (src/raytracer/camera.cpp:120-136) in fp32 for arbitrary yaw/pitch."""
    f32 = np.float32
    yaw, pitch = np.radians(f32(yaw_deg), dtype=f32), np.radians(f32(pitch_deg), dtype=f32)
    front = np.array([np.cos(yaw) * np.cos(pitch), np.sin(pitch), np.sin(yaw) * np.cos(pitch)], f32)
    front = front / np.sqrt((front * front).sum(dtype=f32))
    right = np.cross(front, np.array([0, 1, 0], f32)).astype(f32)
    right = right / np.sqrt((right * right).sum(dtype=f32))
    up = np.cross(right, front).astype(f32)
    up = up / np.sqrt((up * up).sum(dtype=f32))
    return {"origin": tuple(float(v) for v in origin), "direction": tuple(float(v) for v in front),
            "up": tuple(float(v) for v in up), "right": tuple(float(v) for v in right)}


def noise_tables(width, height, seed=0x5EED):
    """(unit[N,3], uniform[N,3]) fp32: unit vectors by rejection from [-1,1]^3 then uniforms in [0,1)
    (same construction as Common::randomUnitVector / randomVec3, include/common/utils.h:30-51)."""
    n = width * height
    rng = np.random.Generator(np.random.PCG64(seed))
    unit = np.empty((n, 3), np.float32)
    filled = 0
    while filled < n:
        p = rng.random((int((n - filled) * 2.2) + 16, 3), dtype=np.float32) * np.float32(2) - np.float32(1)
        l2 = (p * p).sum(axis=1, dtype=np.float32)
        ok = (l2 > 1e-30) & (l2 <= 1)
        p = p[ok] / np.sqrt(l2[ok])[:, None]
        k = min(len(p), n - filled)
        unit[filled:filled + k] = p[:k]
        filled += k
    uniform = rng.random((n, 3), dtype=np.float32)
    return np.ascontiguousarray(unit), np.ascontiguousarray(uniform)


def rubik_scene():
    """Flattened Rubik scene exactly as the reference's loader uploads it (tests/golden/make_scene_golden.py)."""
    z = np.load(os.path.join(GOLDEN, "rubik_scene.npz"))
    return {k: np.ascontiguousarray(z[k]) for k in ("bvhs", "nodes", "materials", "triangles", "vertices")}
