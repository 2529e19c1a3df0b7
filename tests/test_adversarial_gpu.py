"""GPU parity on scenes built to break the hit finder, and on the BASELINE.json configurations at their own size.

The hit finder (csrc/srt_wf_find.cuh) answers rays through a private tree and replays the shader's
box culls; its outside-ray walk rests on a margin argument.  The scenes here aim at the places where such an argument
could fail, each rendered with the finder forced on, forced off, and by the oracle -- all three must agree bit for bit:

  * slivers: large triangles seen edge-on from the camera so that the determinant `a` of the shader's triangle test
    (ray_intersects.glsl:72-74, cut at |a| < 1e-4, then f = 1/a) sits within a factor of a few of the cut;
  * coplanar duplicates and re-triangulated coplanar layers: equal-t ties, decided by visiting order;
  * axis-parallel camera rays whose origin lies ON slab planes of node boxes: 1/0 = inf and 0*inf = NaN inside
    IntersectsBox (ray_intersects.glsl:50-57, Q17);
  * two models whose root boxes overlap (the outside-ray walk is for one BVH; the fallback must hold).

Full size: configs[1] exactly (Rubik 1920x1080x16) and the metric's frame (1080p x 64 spp) with bench.py's own option
set, against the oracle -- the reference's shader text compiled as C++ when oracle/_ref/libshader_ref.so travelled to
this machine, else the hand-written restatement that tests/test_shader_ref.py pins to it.
"""
import os

import numpy as np
import pytest

import scenes
from test_parity_gpu import assert_bit_exact, canon, cpu_render, gpu_render

pytestmark = pytest.mark.gpu


def write_obj(folder, name, verts, groups):
    """groups: [(material name, Ns, Kd, Ks, [(i, j, k) 1-based ...])]"""
    d = os.path.join(folder, name)
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, name + ".mtl"), "w") as f:
        for m, ns, kd, ks, _ in groups:
            f.write("newmtl %s\n\tNs %.4f\n\tKd %.4f %.4f %.4f\n\tKs %.4f %.4f %.4f\n\n" % ((m, ns) + tuple(kd) + tuple(ks)))
    with open(os.path.join(d, name + ".obj"), "w") as f:
        f.write("mtllib %s.mtl\n" % name)
        for v in verts:
            f.write("v %.9g %.9g %.9g\n" % tuple(v))
        for m, _, _, _, faces in groups:
            f.write("usemtl %s\n" % m)
            for fc in faces:
                f.write("f %d %d %d\n" % tuple(fc))
    return name


def grid_wall(verts, faces, x0, x1, y0, y1, z, flip=False):
    """Unit quads with integer corner coordinates in the plane z, two triangles each (the other diagonal if flip)."""
    for x in range(x0, x1):
        for y in range(y0, y1):
            b = len(verts) + 1
            verts.extend([(x, y, z), (x + 1, y, z), (x + 1, y + 1, z), (x, y + 1, z)])
            if flip:
                faces.extend([(b, b + 1, b + 3), (b + 1, b + 2, b + 3)])
            else:
                faces.extend([(b, b + 1, b + 2), (b, b + 2, b + 3)])


def box(verts, faces, lo, hi):
    (x0, y0, z0), (x1, y1, z1) = lo, hi
    b = len(verts) + 1
    verts.extend([(x0, y0, z0), (x1, y0, z0), (x1, y1, z0), (x0, y1, z0), (x0, y0, z1), (x1, y0, z1), (x1, y1, z1), (x0, y1, z1)])
    for q in ((1, 2, 6, 5), (2, 3, 7, 6), (3, 4, 8, 7), (4, 1, 5, 8), (5, 6, 7, 8), (4, 3, 2, 1)):
        a, b_, c, d = (b + i - 1 for i in q)
        faces.extend([(a, b_, c), (a, c, d)])


def sliver_scene(folder, n_slivers=1500, seed=5):
    """A wall + boxes, coplanar duplicates, and triangles whose planes (almost) contain the camera origin."""
    rng = np.random.default_rng(seed)
    C = np.array([0.0, 9.0, 40.0])
    verts, wall, dup, sliv = [], [], [], []
    grid_wall(verts, wall, -8, 8, 1, 17, 0)
    box(verts, wall, (-3, 4, 2), (1, 8, 6))
    box(verts, wall, (2, 10, 1), (5, 13, 4))
    # coplanar layer with the other diagonal over part of the wall, and exact duplicates of every third wall triangle
    grid_wall(verts, dup, -4, 4, 5, 13, 0, flip=True)
    dup.extend(wall[::3])
    for _ in range(n_slivers):
        p1 = np.array([rng.uniform(-9, 9), rng.uniform(0, 18), rng.uniform(1, 14)])
        p2 = p1 + rng.normal(size=3) * rng.uniform(0.5, 3.0)
        n = np.cross(p1 - C, p2 - C)
        n /= np.linalg.norm(n)
        mu = rng.uniform(0.02, 0.12) * rng.choice([-1, 1])
        delta = 10 ** rng.uniform(-5.0, -3.3) * rng.choice([-1, 1])
        p3 = p1 + mu * (p1 - C) + delta * n
        b = len(verts) + 1
        verts.extend([tuple(p1), tuple(p2), tuple(p3)])
        sliv.append((b, b + 1, b + 2))
    groups = [("wall", 10.0, (0.8, 0.7, 0.6), (0.0, 0.0, 0.0), wall), ("dup", 40.0, (0.3, 0.6, 0.9), (0.5, 0.5, 0.5), dup),
              ("sliver", 5.0, (0.9, 0.2, 0.2), (0.2, 0.2, 0.2), sliv)]
    return write_obj(folder, "adversarial", verts, groups)


def three_way(pkg, ctx, oracle, W, H, frames, scene, cam=None, noise=None, max_depth=0):
    """finder forced on (waves to the end), finder off, oracle: all equal."""
    oacc, oout, oids, ocnt = cpu_render(oracle, W, H, True, frames, scene=scene, cam=cam, noise=noise, max_depth=max_depth)
    assert ocnt["capped_paths"] == 0
    res = {}
    for label, sched in (("finder", 31), ("plain", 30), ("defaults", 3)):
        if label == "defaults":
            ctx.set_option(pkg.OPT_FINDER, 1)
            ctx.set_option(pkg.OPT_TAIL_PATHS, 4096)
        acc, out, ids, cnt = gpu_render(pkg, ctx, W, H, True, frames, pkg.TRAV_REFERENCE, scene=scene, cam=cam, noise=noise,
                                        max_depth=max_depth, schedule=sched)
        if label != "plain":
            assert ctx.finder_info()[0], "the finder must be active in this variant"
        assert np.array_equal(ids, oids), label
        assert_bit_exact(acc, oacc)
        assert np.array_equal(out, oout), label
        assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"], label
        res[label] = cnt
    return ocnt


def test_slivers_and_coplanar_duplicates(pkg, ctx, oracle, tmp_path):
    name = sliver_scene(str(tmp_path))
    scene = scenes.load_obj_scene(str(tmp_path), name, str(tmp_path))
    W, H = 512, 288
    ocnt = three_way(pkg, ctx, oracle, W, H, 3, scene)
    assert ocnt["segments"] > W * H * 3 * 1.2  # a sixth of the view is wall: those paths bounce
    # how close to the cut the scene really gets: determinant of the camera rays against the slivers (float64 estimate)
    tri = scene["triangles"]
    v = scene["vertices"]["vertex"].astype(np.float64)
    sl = tri["v"]
    e1 = v[sl[:, 1]] - v[sl[:, 0]]
    e2 = v[sl[:, 2]] - v[sl[:, 0]]
    cen = (v[sl[:, 0]] + v[sl[:, 1]] + v[sl[:, 2]]) / 3.0 - np.array([0.0, 9.0, 40.0])
    d = cen / np.abs(cen[:, 2:3])  # the camera ray through the centroid, |d.z| = 1 as GetRay makes it
    a = np.abs(np.einsum("ij,ij->i", d, np.cross(e1, e2)))
    near = ((a > 0.5e-4) & (a < 2e-4)).sum()
    assert near >= 100, "only %d triangles have |a| within 2x of the 1e-4 cut for their camera ray" % near


def test_axis_parallel_rays_from_slab_planes(pkg, ctx, oracle, tmp_path):
    """Un-jittered camera rays (noise .xy = 0 puts every sample on a pixel corner) from a camera whose basis is exactly
    axis-aligned and whose origin has integer coordinates: column W/2 has direction.x == 0 and row H/2 direction.y == 0
    exactly, and origin.x = 0 / origin.y = 9 are slab planes of node boxes of the integer-grid scene."""
    verts, faces = [], []
    grid_wall(verts, faces, -8, 8, 1, 17, 0)
    box(verts, faces, (0, 9, 2), (4, 13, 6))      # faces ON the planes x = 0 and y = 9
    box(verts, faces, (-5, 3, 1), (0, 9, 5))
    box(verts, faces, (-2, 9, 8), (0, 11, 10))
    name = write_obj(str(tmp_path), "axis", verts, [("m", 10.0, (0.7, 0.7, 0.7), (0.1, 0.1, 0.1), faces)])
    scene = scenes.load_obj_scene(str(tmp_path), name, str(tmp_path))
    assert (scene["nodes"]["min"][:, 0] == 0).any() and (scene["nodes"]["max"][:, 1] == 9).any()
    W, H = 512, 288
    unit, uniform = scenes.noise_tables(W, H)
    unit = unit.copy()
    unit[:, :2] = 0.0
    cam = {"origin": (0.0, 9.0, 40.0), "direction": (0.0, 0.0, -1.0), "up": (0.0, 1.0, 0.0), "right": (1.0, 0.0, 0.0)}
    ocnt = three_way(pkg, ctx, oracle, W, H, 2, scene, cam=cam, noise=(unit, uniform))
    assert ocnt["segments"] > W * H * 2


def test_two_models_with_overlapping_root_boxes(pkg, ctx, oracle, tmp_path):
    """Two interpenetrating copies of Rubik (second one shifted by a third of its size and rotated): every camera ray
    that enters one root box is inside or near the other."""
    from test_parity_gpu import framed
    scene = scenes.concat_scenes(scenes.rubik_scene(), scenes.rubik_scene())
    scene["bvhs"]["frame"][0] = framed(1.0, 0.0, (0.0, 0.0, 0.0))
    scene["bvhs"]["frame"][1] = framed(0.9, 25.0, (5.0, 2.0, 3.0))
    W, H = 384, 216
    oacc, oout, oids, ocnt = cpu_render(oracle, W, H, True, 3, scene=scene)
    n0 = int(scene["bvhs"]["count"][0])
    split = int(scene["nodes"]["index"][n0:][scene["nodes"]["count"][n0:] > 0].min())
    seen = oids[oids != pkg.NO_HIT]
    assert (seen < split).any() and (seen >= split).any()
    for sched in (31, 30, 3):
        if sched == 3:
            ctx.set_option(pkg.OPT_FINDER, 1)
        acc, out, ids, cnt = gpu_render(pkg, ctx, W, H, True, 3, pkg.TRAV_REFERENCE, scene=scene, schedule=sched)
        assert np.array_equal(ids, oids)
        assert_bit_exact(acc, oacc)
        assert np.array_equal(out, oout)
        assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]


# ---- BASELINE.json configurations at their own size, bench.py's option set ------------------------------------------------
def best_oracle(oracle):
    """The reference's shader text compiled as C++ where its .so travelled with the repo, else the hand oracle."""
    return "shader_ref" if oracle.shader_ref_available() else "oracle"


def bench_option_render(pkg, ctx, W, H, spp, scene, noise):
    """What bench.py's timed step does: library defaults + traversal `ordered` (the library walks in the reference order
    wherever the finder answers the rays), one reset dispatch, then spp frames in one srt_render."""
    unit, uniform = noise
    lights, cam = scenes.model_lights(), scenes.reset_camera(True)
    ctx.set_option(pkg.OPT_SCHEDULE, pkg.SCHED_WAVEFRONT)
    ctx.set_option(pkg.OPT_TRAVERSAL, pkg.TRAV_ORDERED)
    ctx.set_option(pkg.OPT_BLOCK_THREADS, 64)
    ctx.upload_noise(unit, uniform)
    ctx.upload_scene(scene)
    ctx.set_lights(lights)
    ctx.resize(W, H)
    ctx.enable_primary_ids(True)
    ctx.set_params(pkg.make_params(W, H, cam, True, accum_frames=1, bvh_count=1, reset=1))
    ctx.render(1)
    ctx.set_params(pkg.make_params(W, H, cam, True, accum_frames=1, bvh_count=1))
    ctx.reset_counters()
    ctx.render(spp)
    return ctx.read_accum(), ctx.read_output(), ctx.read_primary_ids(), ctx.counters()


@pytest.mark.parametrize("spp", [16, 64])
def test_rubik_1080p_against_the_oracle(pkg, ctx, oracle, spp):
    """spp = 16: BASELINE.json configs[1] exactly (objects/Rubik, CPU-built BVH, 1920x1080, 16 spp, 1 B200).
    spp = 64: the frame the metric is quoted on (1080p, 64 spp), as bench.py renders it."""
    W, H = 1920, 1080
    scene = scenes.rubik_scene()
    noise = scenes.noise_tables(W, H)
    acc, out, ids, cnt = bench_option_render(pkg, ctx, W, H, spp, scene, noise)
    assert ctx.finder_info()[0]
    lights, cam = scenes.model_lights(), scenes.reset_camera(True)
    p = oracle.make_params(W, H, cam, True, 1, len(lights), 1, 0)
    oacc, oout, oids, ocnt = oracle.render(p, spp, scene, noise[0], noise[1], lights, impl=best_oracle(oracle))
    assert ocnt["capped_paths"] == 0 and cnt["capped_paths"] == 0
    assert cnt["paths"] == W * H * spp
    assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]
    assert np.array_equal(ids, oids)  # ties budget: 0 pixels
    assert_bit_exact(acc, oacc)
    assert np.array_equal(out, oout)


def test_airplane_standin_1080p_16spp_against_the_oracle(pkg, ctx, oracle, tmp_path):
    """configs[2]'s scene is absent from the reference mount; its documented stand-in at the bench's size and options."""
    W, H, spp = 1920, 1080, 16
    name, _ = scenes.write_airplane_standin(str(tmp_path), detail=128)
    scene = scenes.load_obj_scene(str(tmp_path), name, str(tmp_path))
    noise = scenes.noise_tables(W, H)
    acc, out, ids, cnt = bench_option_render(pkg, ctx, W, H, spp, scene, noise)
    lights, cam = scenes.model_lights(), scenes.reset_camera(True)
    p = oracle.make_params(W, H, cam, True, 1, len(lights), 1, 0)
    oacc, oout, oids, ocnt = oracle.render(p, spp, scene, noise[0], noise[1], lights, impl=best_oracle(oracle))
    assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]
    bad = (canon(acc) != canon(oacc)).any(axis=2)
    # bench.py passes traversal=ordered; where the finder is active the library walks in the reference order and the
    # budget below is never used (measured: 0 pixels), where it is not, documented grazing ties are allowed
    budget = 0 if ctx.finder_info()[0] else 1e-4
    assert bad.mean() <= budget, "%d pixels differ" % bad.sum()
    assert (ids != oids).mean() <= budget
