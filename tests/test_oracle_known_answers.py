"""Known-answer tests of the oracle's building blocks against independent numpy restatements (CPU only).

These do not need bit equality (numpy has no fma): they check that the transcription computes the right THING --
the decree's sin/cos/pow against float64, the sRGB curve of raytrace_utils.glsl:177-184, the BVH traversal of
ray_intersects.glsl:99-133 against a brute-force Moller-Trumbore over all triangles, and the primary hits of the
built-in sphere scene (raytrace_compute.glsl:299-364) against the analytic ray/sphere solution.
"""
import numpy as np

import scenes


def ulp_diff(a, b):
    a = np.asarray(a, np.float32)
    b = np.asarray(b, np.float32)
    ia = a.view(np.int32).astype(np.int64)
    ib = b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7FFFFFFF), ia)
    ib = np.where(ib < 0, -(ib & 0x7FFFFFFF), ib)
    return np.abs(ia - ib)


def test_decree_sin_cos_close_to_float64(oracle):
    L = oracle.lib()
    rng = np.random.default_rng(7)
    x = np.concatenate([rng.uniform(-200, 200, 4000), rng.uniform(-9000, 9000, 4000), rng.uniform(-1e-3, 1e-3, 500)])
    x = x.astype(np.float32)
    s = np.array([L.oracle_sin(float(v)) for v in x], np.float32)
    c = np.array([L.oracle_cos(float(v)) for v in x], np.float32)
    # absolute error: the seeds of randFloat are O(10^3), where 1 ulp of the argument is already 6e-5 of a period
    assert np.abs(s - np.sin(x.astype(np.float64))).max() < 4e-7
    assert np.abs(c - np.cos(x.astype(np.float64))).max() < 4e-7
    assert L.oracle_sin(0.0) == 0.0 and L.oracle_cos(0.0) == 1.0
    assert np.isnan(L.oracle_sin(float("inf"))) and np.isnan(L.oracle_cos(float("nan")))


def test_decree_pow_and_srgb_curve(oracle):
    L = oracle.lib()
    rng = np.random.default_rng(8)
    x = rng.uniform(0.0, 4.0, 3000).astype(np.float32)
    got = np.array([L.oracle_pow(float(v), float(np.float32(1.0 / 2.4))) for v in x], np.float32)
    want = np.power(x.astype(np.float64), float(np.float32(1.0 / 2.4)))
    assert (np.abs(got - want) <= 4e-7 * np.maximum(want, 1e-3)).all()
    srgb = np.array([L.oracle_linear_to_srgb(float(v)) for v in x], np.float32)
    xd = x.astype(np.float64)
    ref = np.where(xd <= 0.0031308, 12.92 * xd, 1.055 * np.power(xd, 1 / 2.4) - 0.055)
    assert np.abs(srgb - ref).max() < 2e-6 * 4
    assert L.oracle_linear_to_srgb(0.0) == 0.0


def test_rand_float_is_fract_of_scaled_sine(oracle):
    L = oracle.lib()
    rng = np.random.default_rng(9)
    for sx, sy in rng.uniform(-30, 30, (2000, 2)).astype(np.float32):
        r = L.oracle_rand_float(float(sx), float(sy))
        assert 0.0 <= r <= 1.0  # fract() of an fp32 can round up to exactly 1.0
        # dot(seed, k) as the decree contracts it: fma(sy, k.y, fl(sx * k.x)); float64 holds both products exactly
        first = np.float64(np.float32(np.float32(sx) * np.float32(12.9898)))
        d = np.float32(np.float64(sy) * np.float64(np.float32(78.233)) + first)
        want = (np.sin(np.float64(d)) * np.float64(np.float32(43758.5453))) % 1.0
        # sin(d)*43758.5453 amplifies the decree's 1e-7 absolute error by 4e4: agreement to ~1e-2 is all that is defined
        assert min(abs(r - want), 1 - abs(r - want)) < 3e-2


def brute_force(scene, o, d, tmax):
    tri = scene["triangles"]["v"]
    v = scene["vertices"]["vertex"].astype(np.float64)
    v0, v1, v2 = v[tri[:, 0]], v[tri[:, 1]], v[tri[:, 2]]
    e1, e2 = v1 - v0, v2 - v0
    h = np.cross(d, e2)
    a = (e1 * h).sum(1)
    ok = ~((a > -1e-4) & (a < 1e-4))
    f = 1.0 / np.where(ok, a, 1.0)
    s = o - v0
    u = f * (s * h).sum(1)
    q = np.cross(s, e1)
    w = f * (q @ d)
    t = f * (e2 * q).sum(1)
    ok &= (u >= 0) & (u <= 1) & (w >= 0) & (u + w <= 1) & (t > 1e-5) & (t < tmax)
    t = np.where(ok, t, np.inf)
    return t


def test_bvh_traversal_matches_brute_force_from_outside(oracle):
    """Rays that start outside the root box (so no box ever contains the origin and the shader's exit-distance cull,
    ray_intersects.glsl:57, cannot drop a subtree): closest hit == brute force over all 1188 triangles."""
    scene = scenes.rubik_scene()
    rng = np.random.default_rng(11)
    n = 600
    lo, hi = scene["nodes"]["min"][0], scene["nodes"]["max"][0]
    centre, radius = (lo + hi) / 2, np.linalg.norm(hi - lo)
    dirs = rng.normal(size=(n, 3))
    dirs /= np.linalg.norm(dirs, axis=1)[:, None]
    origins = centre + dirs * radius * 1.5
    targets = rng.uniform(lo, hi, (n, 3))
    rays = np.zeros(n, oracle.RAY_DT)
    rays["origin"] = origins.astype(np.float32)
    rays["direction"] = (targets - origins).astype(np.float32)
    rays["dist"] = 1e30
    hit, t, cnt = oracle.trace_rays(scene, rays)
    n_hits = 0
    for i in range(n):
        tt = brute_force(scene, rays["origin"][i].astype(np.float64), rays["direction"][i].astype(np.float64), 1e30)
        best = tt.min()
        if np.isinf(best):
            assert hit[i] == 0xFFFFFFFF, i
            continue
        n_hits += 1
        assert hit[i] != 0xFFFFFFFF, i
        assert abs(t[i] - best) <= 2e-5 * best, (i, t[i], best)
        assert tt[hit[i]] <= best * (1 + 2e-5)  # the winner is (one of) the closest
    assert n_hits > n // 3
    assert cnt["nodes"] > 0 and cnt["tris"] > 0


def test_sphere_scene_primary_hits_match_analytic_solution(oracle):
    W, H = 96, 72
    unit, uniform = scenes.noise_tables(W, H)
    unit = unit.copy()
    unit[:, :2] = 0.5  # SampleSquare = noise.xy - 0.5 = 0: rays through pixel centres
    lights, cam = scenes.sphere_lights(), scenes.reset_camera(False)
    p = oracle.make_params(W, H, cam, False, 1, len(lights), 0)
    _, _, ids, _ = oracle.render(p, 1, None, unit, uniform, lights)
    spheres = np.array([(1.8, 0.0, -2.0, 0.5), (0.0, -100.5, -1.0, 100.0), (0.55, 0.0, -2.0, 0.5),
                        (-0.55, 0.0, -2.0, 0.5), (-1.8, 0.0, -2.0, 0.5)])  # raytrace_compute.glsl:340-364, world[0..4]
    o = np.array(cam["origin"])
    w = -np.array(cam["direction"], np.float64)
    u, v = np.array(cam["right"], np.float64), np.array(cam["up"], np.float64)
    du, dv = u / W, v / H  # 1x1 viewport at distance 1 (Q12)
    p00 = o - w - u / 2 - v / 2 + 0.5 * (du + dv)
    mism = 0
    for y in range(H):
        for x in range(W):
            d = p00 + x * du + y * dv - o
            best, win = np.inf, 0xFFFFFFFF
            for k, (cx, cy, cz, r) in enumerate(spheres):
                oc = np.array([cx, cy, cz]) - o
                a, hh, c = d @ d, d @ oc, oc @ oc - r * r
                disc = hh * hh - a * c
                if disc < 0:
                    continue
                for root in ((hh - np.sqrt(disc)) / a, (hh + np.sqrt(disc)) / a):
                    if 0.001 < root < best:
                        best, win = root, k
                        break
            mism += int(win != ids[y, x])
    assert mism <= 3, mism  # silhouette pixels may flip between fp32 and fp64
