"""GPU parity tests: the CUDA path through the C-ABI against the CPU oracle on the same seeded inputs.

Bar (BASELINE.json north_star): primary-hit ids bit-exact, radiance within 1e-4 relative.  With the
reference visiting order (SRT_OPT_TRAVERSAL=0) the kernels share every expression tree with the oracle, so
the tests below demand BIT-EXACT fp32 accumulation images; the ordered traversal is held to "identical except
documented grazing ties", with the tie budget written in the test.
"""
import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu


def gpu_render(pkg, ctx, W, H, show_model, n_frames, traversal, scene=None, lights=None, cam=None, first_frame=1,
               max_depth=0, noise=None, rows=None, schedule=3, textures=None, tail=None, finder=None):
    """Variants of the wavefront schedule (3): plain 3 = the library's defaults (hit finder where srt_upload_scene finds
    it worthwhile -- Rubik yes, soup no --, and launches of these sizes run in the tail kernel right after
    wf_init_kernel); 30 = waves to the very end (SRT_OPT_TAIL_PATHS = 0) walking the reference BVH (SRT_OPT_FINDER = 0);
    31 = waves to the very end through the hit finder (SRT_OPT_FINDER = 1); 32 = default tail, finder off; 33 = the
    defaults with SRT_OPT_PRIMARY_CULL = 0 (every primary ray through the queues; the others end root-box misses in
    wf_init_kernel)."""
    if schedule == 30:
        schedule, tail, finder = 3, 0, 0
    elif schedule == 31:
        schedule, tail, finder = 3, 0, 1
    elif schedule == 32:
        schedule, finder = 3, 0
    elif schedule == 33:
        schedule = 3
        ctx.set_option(pkg.OPT_PRIMARY_CULL, 0)
    if tail is not None:
        ctx.set_option(pkg.OPT_TAIL_PATHS, tail)
    if finder is not None:
        ctx.set_option(pkg.OPT_FINDER, finder)
    unit, uniform = noise if noise is not None else scenes.noise_tables(W, H)
    lights = lights if lights is not None else (scenes.model_lights() if show_model else scenes.sphere_lights())
    cam = cam or scenes.reset_camera(show_model)
    ctx.set_option(pkg.OPT_TRAVERSAL, traversal)
    ctx.set_option(pkg.OPT_SCHEDULE, schedule)
    if scene is not None:
        ctx.upload_scene(scene)
        for material, image in textures or []:
            ctx.set_texture(material, image)
    ctx.upload_noise(unit, uniform)
    ctx.set_lights(lights)
    ctx.resize(W, H)
    ctx.enable_primary_ids(True)
    if rows:
        ctx.set_rows(*rows)
    ctx.set_params(pkg.make_params(W, H, cam, show_model, accum_frames=first_frame, max_depth=max_depth,
                                   bvh_count=len(scene["bvhs"]) if scene is not None else 0))
    ctx.reset_counters()
    ctx.render(n_frames)
    return ctx.read_accum(), ctx.read_output(), ctx.read_primary_ids(), ctx.counters()


def cpu_render(oracle, W, H, show_model, n_frames, scene=None, lights=None, cam=None, first_frame=1, max_depth=0,
               noise=None):
    unit, uniform = noise if noise is not None else scenes.noise_tables(W, H)
    lights = lights if lights is not None else (scenes.model_lights() if show_model else scenes.sphere_lights())
    cam = cam or scenes.reset_camera(show_model)
    p = oracle.make_params(W, H, cam, show_model, first_frame, len(lights), len(scene["bvhs"]) if scene else 0, max_depth)
    return oracle.render(p, n_frames, scene, unit, uniform, lights)


def canon(a):
    """fp32 bit patterns with every NaN mapped to one pattern (the payload of a NaN is not part of the result)."""
    a = np.ascontiguousarray(a, np.float32)
    return np.where(np.isnan(a), np.uint32(0x7FC00000), a.view(np.uint32))


def assert_bit_exact(acc, oacc):
    a, b = canon(acc), canon(oacc)
    if not np.array_equal(a, b):
        bad = (a != b).any(axis=2)
        ys, xs = np.nonzero(bad)
        raise AssertionError("%d of %d pixels differ; first (x=%d,y=%d): gpu %r oracle %r" % (
            bad.sum(), bad.size, xs[0], ys[0], acc[ys[0], xs[0]], oacc[ys[0], xs[0]]))


def test_detmath_bitwise(pkg, ctx, oracle):
    rng = np.random.default_rng(1)
    x = np.concatenate([rng.uniform(-8000, 8000, 200000), rng.uniform(-1, 1, 50000) * 1e-3,
                        [0.0, -0.0, np.inf, -np.inf, np.nan, 1e9, -3e8, 1e-40]]).astype(np.float32)
    L = oracle.lib()
    for which, fn in ((0, L.oracle_sin), (1, L.oracle_cos)):
        got = ctx.eval_detmath(which, x)
        ref = np.array([fn(float(v)) for v in x[:20000].tolist() + x[-8:].tolist()], np.float32)
        g = np.concatenate([got[:20000], got[-8:]])
        assert np.array_equal(canon(g), canon(ref))
    y = rng.uniform(-50, 50, x.size).astype(np.float32)
    got = ctx.eval_detmath(3, x[:20000], y[:20000])
    ref = np.array([L.oracle_rand_float(float(a), float(b)) for a, b in zip(x[:20000], y[:20000])], np.float32)
    assert np.array_equal(canon(got), canon(ref))
    c = np.concatenate([rng.uniform(0, 4, 20000), [0.0, 0.0031308, 1.0, np.inf, np.nan, -1.0]]).astype(np.float32)
    got = ctx.eval_detmath(4, c)
    ref = np.array([L.oracle_linear_to_srgb(float(v)) for v in c], np.float32)
    assert np.array_equal(canon(got), canon(ref))


@pytest.mark.parametrize("schedule", [0, 1])
def test_c1_spheres_640x480_1spp_bit_exact(pkg, ctx, oracle, schedule):
    """BASELINE.json configs[0]: built-in spheres, 640x480, 1 spp."""
    acc, out, ids, cnt = gpu_render(pkg, ctx, 640, 480, False, 1, pkg.TRAV_REFERENCE, schedule=schedule)
    oacc, oout, oids, ocnt = cpu_render(oracle, 640, 480, False, 1)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert np.array_equal(out, oout)
    assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]
    assert cnt["paths"] == 640 * 480


@pytest.mark.parametrize("frames", [(1, 3), (7, 2)])
def test_spheres_multi_frame_bit_exact(pkg, ctx, oracle, frames):
    first, n = frames
    acc, out, ids, _ = gpu_render(pkg, ctx, 320, 240, False, n, pkg.TRAV_REFERENCE, first_frame=first)
    oacc, oout, oids, _ = cpu_render(oracle, 320, 240, False, n, first_frame=first)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert np.array_equal(out, oout)


@pytest.mark.parametrize("schedule", [0, 1, 2, 3, 30, 31, 32, 33])
def test_rubik_reference_order_bit_exact(pkg, ctx, oracle, schedule):
    """C2's scene at a size the oracle finishes in seconds; reference visiting order => bit-exact."""
    scene = scenes.rubik_scene()
    acc, out, ids, cnt = gpu_render(pkg, ctx, 640, 360, True, 4, pkg.TRAV_REFERENCE, scene=scene, schedule=schedule)
    oacc, oout, oids, ocnt = cpu_render(oracle, 640, 360, True, 4, scene=scene)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert np.array_equal(out, oout)
    assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]
    assert ocnt["capped_paths"] == 0 and cnt["capped_paths"] == 0


@pytest.mark.parametrize("schedule", [0, 1, 2, 3, 30, 31, 32])
def test_rubik_ordered_traversal_matches_bar_ties(pkg, ctx, oracle, schedule):
    """Ordered traversal: same image except documented grazing ties (budget: 1e-4 of the pixels; ids identical
    on every pixel outside that budget, radiance bit-exact there too)."""
    scene = scenes.rubik_scene()
    acc, out, ids, cnt = gpu_render(pkg, ctx, 640, 360, True, 4, pkg.TRAV_ORDERED, scene=scene, schedule=schedule)
    oacc, oout, oids, ocnt = cpu_render(oracle, 640, 360, True, 4, scene=scene)
    bad = (acc.view(np.uint32) != oacc.view(np.uint32)).any(axis=2)
    assert bad.mean() <= 1e-4, "ordered traversal differs on %d pixels" % bad.sum()
    assert (ids != oids).mean() <= 1e-4


def test_wavefront_small_pool_and_frame_chunks_bit_exact(pkg, ctx, oracle):
    """Wavefront schedule with a pool far smaller than the launch (slots regenerate many times) and a sample budget
    that cuts the launch into several frame chunks: still the oracle's bits."""
    scene = scenes.rubik_scene()
    ctx.set_option(pkg.OPT_WAVE_SLOTS, 4096)
    ctx.set_option(pkg.OPT_SAMPLE_BUDGET_MB, 1)  # 1 MB / 16 B = 65536 entries = 2 frames of 160x96... per chunk
    acc, out, ids, cnt = gpu_render(pkg, ctx, 256, 128, True, 5, pkg.TRAV_REFERENCE, scene=scene, schedule=3)
    oacc, oout, oids, ocnt = cpu_render(oracle, 256, 128, True, 5, scene=scene)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert np.array_equal(out, oout)
    assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]
    assert cnt["paths"] == 256 * 128 * 5


@pytest.mark.parametrize("schedule", [2, 3, 30, 31])
def test_airplane_standin_bit_exact(pkg, ctx, oracle, tmp_path, schedule):
    """BASELINE.json configs[2]'s scene is absent from the reference mount; its documented stand-in (6 Airplane
    materials, leaves of up to ~100 triangles from the reference's midpoint builder) through the host loader."""
    name, n_tri = scenes.write_airplane_standin(str(tmp_path), detail=40)
    scene = scenes.load_obj_scene(str(tmp_path), name, str(tmp_path))
    assert len(scene["triangles"]) == n_tri and len(scene["materials"]) == 6
    acc, out, ids, cnt = gpu_render(pkg, ctx, 320, 184, True, 3, pkg.TRAV_REFERENCE, scene=scene, schedule=schedule)
    oacc, oout, oids, ocnt = cpu_render(oracle, 320, 184, True, 3, scene=scene)
    assert (oids != pkg.NO_HIT).mean() > 0.05
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert np.array_equal(out, oout)
    assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]
    acc2, _, ids2, _ = gpu_render(pkg, ctx, 320, 184, True, 3, pkg.TRAV_ORDERED, scene=scene, schedule=schedule)
    bad = (canon(acc2) != canon(oacc)).any(axis=2)
    assert bad.mean() <= 1e-4 and (ids2 != oids).mean() <= 1e-4  # documented grazing ties only


def framed(scale, rot_y_deg, translate):
    """World -> model matrix of a model placed with (scale, rotation about y, translation), column-major like
    BVH.frame (raytrace_types.glsl:37): the inverse of T * R * S."""
    c, s_ = np.cos(np.radians(rot_y_deg)), np.sin(np.radians(rot_y_deg))
    R = np.array([[c, 0, s_, 0], [0, 1, 0, 0], [-s_, 0, c, 0], [0, 0, 0, 1]], np.float64)
    S = np.diag([scale, scale, scale, 1.0])
    T = np.eye(4)
    T[:3, 3] = translate
    return np.linalg.inv(T @ R @ S).astype(np.float32).T.reshape(16).copy()


@pytest.mark.parametrize("schedule", [0, 2, 3, 30, 31, 33])
def test_two_models_with_frames_bit_exact(pkg, ctx, oracle, tmp_path, schedule):
    """bvh_count = 2, each model behind its own world->model frame (raytrace_compute.glsl:144-161);
    the closest hit is taken over both, t shared between model spaces as the shader does (Q14).  Then one frame is
    replaced through UpdateModelMatrix (gpu_loader.cpp:185-196) and the picture must follow."""
    name, _ = scenes.write_airplane_standin(str(tmp_path), detail=24)
    scene = scenes.concat_scenes(scenes.rubik_scene(), scenes.load_obj_scene(str(tmp_path), name, str(tmp_path)))
    assert len(scene["bvhs"]) == 2
    scene["bvhs"]["frame"][0] = framed(0.6, 30.0, (-7.0, 2.0, 0.0))
    scene["bvhs"]["frame"][1] = framed(1.3, -20.0, (6.0, 1.0, 5.0))
    for mode in (pkg.TRAV_REFERENCE,):
        acc, out, ids, cnt = gpu_render(pkg, ctx, 256, 128, True, 3, mode, scene=scene, schedule=schedule)
        oacc, oout, oids, ocnt = cpu_render(oracle, 256, 128, True, 3, scene=scene)
        first = int(scene["bvhs"]["count"][0])
        tri_split = int(scene["nodes"]["index"][first:][scene["nodes"]["count"][first:] > 0].min())
        seen = oids[oids != pkg.NO_HIT]
        assert (seen < tri_split).any() and (seen >= tri_split).any()  # both models are in the picture
        assert np.array_equal(ids, oids)
        assert_bit_exact(acc, oacc)
        assert np.array_equal(out, oout)
        assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]
    moved = framed(1.3, 75.0, (2.0, 4.0, -3.0))
    ctx.update_frame(1, moved)
    ctx.resize(256, 128)  # fresh (zeroed) accumulation image
    ctx.enable_primary_ids(True)
    ctx.set_params(pkg.make_params(256, 128, scenes.reset_camera(True), True, accum_frames=1, bvh_count=2))
    ctx.render(2)
    scene["bvhs"]["frame"][1] = moved
    oacc2, _, oids2, _ = cpu_render(oracle, 256, 128, True, 2, scene=scene)
    assert np.array_equal(ctx.read_primary_ids(), oids2) and not np.array_equal(oids2, oids)
    assert_bit_exact(ctx.read_accum(), oacc2)


@pytest.mark.parametrize("schedule", [3, 30, 31])
def test_soup_depth8_bit_exact(pkg, ctx, oracle, tmp_path, schedule):
    """BASELINE.json configs[4] at a size the oracle finishes in seconds: random soup, maxDepth 8, camera (0,0,120)."""
    scene, info = scenes.soup_scene(20000, str(tmp_path))
    assert info["depth"] <= 63
    cam = scenes.reset_camera(True, origin=(0.0, 0.0, 120.0))
    acc, out, ids, cnt = gpu_render(pkg, ctx, 256, 144, True, 2, pkg.TRAV_REFERENCE, scene=scene, cam=cam, max_depth=8,
                                    schedule=schedule)
    oacc, oout, oids, ocnt = cpu_render(oracle, 256, 144, True, 2, scene=scene, cam=cam, max_depth=8)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]
    assert ocnt["capped_paths"] == 0


def test_one_trace_launch_per_queue_changes_no_bit(pkg, ctx, oracle, tmp_path, monkeypatch):
    """Scenes far larger than L2 are traced queue by queue (closest-hit rays, then shadow rays, a launch each with its own
    cursor); SRT_QUEUE_BY_QUEUE forces that at a size the oracle finishes.  Waves to the very end on the reference BVH."""
    monkeypatch.setenv("SRT_QUEUE_BY_QUEUE", "1")
    scene, info = scenes.soup_scene(20000, str(tmp_path))
    cam = scenes.reset_camera(True, origin=(0.0, 0.0, 120.0))
    for schedule in (30, 32):
        acc, out, ids, cnt = gpu_render(pkg, ctx, 256, 144, True, 2, pkg.TRAV_REFERENCE, scene=scene, cam=cam, max_depth=8,
                                        schedule=schedule)
        oacc, oout, oids, ocnt = cpu_render(oracle, 256, 144, True, 2, scene=scene, cam=cam, max_depth=8)
        assert np.array_equal(ids, oids)
        assert_bit_exact(acc, oacc)
        assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]


@pytest.mark.parametrize("schedule", [3, 30, 31])
def test_rubik_inside_view_and_depth8(pkg, ctx, oracle, schedule):
    """Camera inside the model's bounds looking around (many bounces), maxDepth 8 as C5 uses."""
    scene = scenes.rubik_scene()
    cam = scenes.look_camera((0.3, 9.1, 12.0), -70.0, -12.0)
    acc, out, ids, _ = gpu_render(pkg, ctx, 256, 256, True, 3, pkg.TRAV_REFERENCE, scene=scene, cam=cam, max_depth=8,
                                  schedule=schedule)
    oacc, oout, oids, _ = cpu_render(oracle, 256, 256, True, 3, scene=scene, cam=cam, max_depth=8)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)


def test_primary_cull_same_bits_same_counts(pkg, ctx, oracle):
    """SRT_OPT_PRIMARY_CULL: primary rays that miss every root box end in wf_init_kernel.  Same picture, same segment
    and shadow-ray counts, and the counting pass (nodes box-tested, triangles tested in the shader's visiting order)
    reports the same totals as with every primary ray walked by the trace kernel; a view that sees no model at all
    (every path culled, wave 0 empty) is the sky colour everywhere."""
    scene = scenes.rubik_scene()
    res = {}
    for cull in (1, 0):
        ctx.set_option(pkg.OPT_PRIMARY_CULL, cull)
        ctx.set_option(pkg.OPT_COUNT_TRAVERSAL, 1)
        res[cull] = gpu_render(pkg, ctx, 320, 184, True, 3, pkg.TRAV_REFERENCE, scene=scene, schedule=30)
    ctx.set_option(pkg.OPT_COUNT_TRAVERSAL, 0)
    oacc, oout, oids, ocnt = cpu_render(oracle, 320, 184, True, 3, scene=scene)
    for cull in (1, 0):
        acc, out, ids, cnt = res[cull]
        assert np.array_equal(ids, oids)
        assert_bit_exact(acc, oacc)
        assert np.array_equal(out, oout)
        assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]
    assert res[1][3]["nodes"] == res[0][3]["nodes"] and res[1][3]["tris"] == res[0][3]["tris"]
    assert res[1][3]["nodes"] > 320 * 184 * 3
    # looking away from the model: nothing survives the cull
    ctx.set_option(pkg.OPT_PRIMARY_CULL, 1)
    cam = scenes.look_camera((0.0, 9.0, 40.0), 90.0, 0.0)
    acc, out, ids, cnt = gpu_render(pkg, ctx, 320, 184, True, 2, pkg.TRAV_REFERENCE, scene=scene, cam=cam, schedule=3)
    oacc, oout, oids, ocnt = cpu_render(oracle, 320, 184, True, 2, scene=scene, cam=cam)
    assert (oids == pkg.NO_HIT).all() and np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert np.array_equal(out, oout)
    assert cnt["segments"] == ocnt["segments"] == 320 * 184 * 2 and cnt["shadow_rays"] == 0


def test_pinned_noise_tables_same_picture(pkg, ctx, oracle):
    """srt_upload_noise takes page-locked tables by DMA and packs them on the device (pageable ones are packed by host
    threads): the same picture either way, and a scene re-upload into the kept device allocations changes nothing."""
    import torch
    scene = scenes.rubik_scene()
    W, H = 320, 184
    unit, uniform = scenes.noise_tables(W, H)
    keep = [torch.from_numpy(np.ascontiguousarray(a, np.float32)).pin_memory() for a in (unit, uniform)]
    acc, out, ids, _ = gpu_render(pkg, ctx, W, H, True, 2, pkg.TRAV_REFERENCE, scene=scene, noise=(keep[0].numpy(), keep[1].numpy()))
    oacc, oout, oids, _ = cpu_render(oracle, W, H, True, 2, scene=scene)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    # second upload of the same scene (device arrays reused), pageable tables this time
    acc2, _, ids2, _ = gpu_render(pkg, ctx, W, H, True, 2, pkg.TRAV_REFERENCE, scene=scene, noise=(unit, uniform))
    assert np.array_equal(ids2, oids)
    assert_bit_exact(acc2, oacc)


@pytest.mark.parametrize("which", ["rubik", "airplane-standin"])
def test_full_size_defaults_equal_the_plain_walk(pkg, ctx, tmp_path, which):
    """At the bench's full size (1920x1080, 64 spp in one launch: 132.7 M paths, too many for the oracle) the library's
    defaults -- hit finder with its distance-bounded walk for camera rays, primary cull, tail kernel, no-light paths put
    aside -- must give the very bits of the plain schedule that the tests above pin to the oracle at smaller sizes: every
    ray through wf_trace_kernel in the shader's visiting order, waves to the end, every camera ray through the queues."""
    if which == "rubik":
        scene = scenes.rubik_scene()
    else:
        name, _ = scenes.write_airplane_standin(str(tmp_path), detail=128)
        scene = scenes.load_obj_scene(str(tmp_path), name, str(tmp_path))
    W, H, spp = 1920, 1080, 64
    noise = scenes.noise_tables(W, H)
    ctx.set_option(pkg.OPT_FINDER, 1)
    acc, out, ids, cnt = gpu_render(pkg, ctx, W, H, True, spp, pkg.TRAV_REFERENCE, scene=scene, noise=noise, schedule=3)
    assert ctx.finder_info()[0]
    ctx.set_option(pkg.OPT_PRIMARY_CULL, 0)
    acc0, out0, ids0, cnt0 = gpu_render(pkg, ctx, W, H, True, spp, pkg.TRAV_REFERENCE, scene=scene, noise=noise, schedule=30)
    assert not ctx.finder_info()[0]
    assert cnt["paths"] == W * H * spp == cnt0["paths"]
    assert cnt["segments"] == cnt0["segments"] and cnt["shadow_rays"] == cnt0["shadow_rays"]
    assert np.array_equal(ids, ids0)
    assert np.array_equal(acc.view(np.uint32), acc0.view(np.uint32))
    assert np.array_equal(out, out0)


def test_upload_rejects_bad_indices_with_the_lowest_offender(pkg, ctx, tmp_path):
    """srt_upload_scene validates every triangle (the records are built on host threads for big scenes): the error names
    the lowest offending triangle whatever the thread count, a vertex index before a material index on the same one."""
    scene, _ = scenes.soup_scene(70000, str(tmp_path))  # above the threading threshold (64 Ki triangles)
    n_vert, n_mat = len(scene["vertices"]), len(scene["materials"])
    bad = {k: v.copy() for k, v in scene.items() if hasattr(v, "copy")}
    bad["triangles"]["v"][69000, 1] = n_vert
    bad["triangles"]["v"][1234, 2] = n_vert + 5
    bad["triangles"]["material"][1234] = n_mat
    bad["triangles"]["material"][77] = n_mat + 1
    with pytest.raises(pkg.SrtError, match="triangle 77: material index out of range"):
        ctx.upload_scene(bad)
    bad["triangles"]["material"][77] = 0
    with pytest.raises(pkg.SrtError, match="triangle 1234: vertex index out of range"):
        ctx.upload_scene(bad)
    ctx.upload_scene(scene)  # and a good scene still goes in afterwards


def test_row_bands_equal_whole_image(pkg, ctx, oracle):
    """The image-tile half of the multi-GPU split: rendering two row bands gives the bits of one full render."""
    scene = scenes.rubik_scene()
    W, H = 320, 256
    whole, _, _, _ = gpu_render(pkg, ctx, W, H, True, 2, pkg.TRAV_REFERENCE, scene=scene)
    top, _, _, _ = gpu_render(pkg, ctx, W, H, True, 2, pkg.TRAV_REFERENCE, scene=scene, rows=(0, 96))
    ctx.set_rows(96, H)
    ctx.set_params(pkg.make_params(W, H, scenes.reset_camera(True), True, accum_frames=1, bvh_count=1))
    ctx.render(2)
    both = ctx.read_accum()
    assert np.array_equal(both.view(np.uint32), whole.view(np.uint32))


@pytest.mark.parametrize("world", [2, 3])
def test_interleaved_stripes_sum_to_whole_image(pkg, ctx, oracle, world):
    """The multi-GPU split on one device: every rank's context renders its interleaved 8-row bands (srt_set_stripe),
    rows it does not own stay +0.0, and the SUM of the ranks' accumulation images (what the NCCL reduce computes)
    has the bits of the whole-image render and of the oracle."""
    scene = scenes.rubik_scene()
    W, H, frames = 320, 200, 3
    whole, _, _, cnt = gpu_render(pkg, ctx, W, H, True, frames, pkg.TRAV_REFERENCE, scene=scene)
    total = np.zeros_like(whole)
    segs = 0
    for rank in range(world):
        c = pkg.Context(0)
        try:
            c.set_stripe(rank, world)
            part, _, _, pc = gpu_render(pkg, c, W, H, True, frames, pkg.TRAV_REFERENCE, scene=scene)
        finally:
            c.close()
        own = np.array([(y // 8) % world == rank for y in range(H)])
        assert not part[~own][..., :3].any()  # untouched rows are exactly zero
        total = total + part
        segs += pc["segments"]
    assert np.array_equal(canon(total[..., :3]), canon(whole[..., :3]))
    assert segs == cnt["segments"]
    oacc, _, _, _ = cpu_render(oracle, W, H, True, frames, scene=scene)
    assert np.array_equal(canon(total[..., :3]), canon(oacc[..., :3]))


def test_trace_rays_reference_fixture_shape(pkg, ctx, oracle):
    """The reference's (disabled, stale) 64-ray test re-cast with oracle-derived expectations
    (include/compute/tests/BVH_intergration_tests.cpp:69-113)."""
    scene = scenes.rubik_scene()
    rays = np.zeros(64, pkg.RAY_DT)
    for i in range(64):
        if i % 2:
            rays[i]["origin"], rays[i]["direction"] = (-10.0, 3.0, 6.0), (0.9838, -0.0118, 0.1787)
        else:
            rays[i]["origin"], rays[i]["direction"] = (0.0, 0.0, 0.0), (0.0, 1.0, 0.0)
        rays[i]["dist"] = 1e30
    ohit, ot, _ = oracle.trace_rays(scene, rays)
    for mode in (pkg.TRAV_REFERENCE, pkg.TRAV_ORDERED):
        ctx.set_option(pkg.OPT_TRAVERSAL, mode)
        ctx.upload_scene(scene)
        hit, t = ctx.trace_rays(rays)
        assert np.array_equal(hit, ohit) and np.array_equal(t.view(np.uint32), ot.view(np.uint32))
    # second half of the reference test: a degenerate frame makes every ray miss
    m = np.full((4, 4), 0.0, np.float32)
    np.fill_diagonal(m, 1e-6)
    m[0, 3], m[1, 3], m[2, 3] = 10, 1000, 10  # glm new_mat[c][r]
    ctx.update_frame(0, m.reshape(16))
    hit, _ = ctx.trace_rays(rays)
    scene2 = {k: v.copy() for k, v in scene.items()}
    scene2["bvhs"]["frame"][0] = m.reshape(16)
    ohit2, _, _ = oracle.trace_rays(scene2, rays)
    assert np.array_equal(hit, ohit2)
    assert (hit == pkg.NO_HIT).all()


@pytest.mark.parametrize("schedule", [3, 30, 31])
def test_sorted_ray_queues_change_no_bit(pkg, ctx, oracle, tmp_path, schedule):
    """SRT_OPT_SORT_RAYS: the wave's queues regrouped by the cell the rays start in (counting sort by a 15-bit Morton code,
    csrc/srt_wf_sort.cuh) -- another ORDER of answering independent rays, the same picture: soup at depth 8 (the case it is
    for), Rubik from inside, and a small pool whose slots regenerate (holes in the queues)."""
    ctx.set_option(pkg.OPT_SORT_RAYS, 1)
    scene, _ = scenes.soup_scene(20000, str(tmp_path))
    cam = scenes.reset_camera(True, origin=(0.0, 0.0, 120.0))
    acc, out, ids, cnt = gpu_render(pkg, ctx, 256, 144, True, 2, pkg.TRAV_REFERENCE, scene=scene, cam=cam, max_depth=8,
                                    schedule=schedule)
    oacc, oout, oids, ocnt = cpu_render(oracle, 256, 144, True, 2, scene=scene, cam=cam, max_depth=8)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]
    rubik = scenes.rubik_scene()
    inside = scenes.look_camera((0.3, 9.1, 12.0), -70.0, -12.0)
    ctx.set_option(pkg.OPT_WAVE_SLOTS, 8192)
    acc, out, ids, cnt = gpu_render(pkg, ctx, 256, 256, True, 3, pkg.TRAV_REFERENCE, scene=rubik, cam=inside, max_depth=8,
                                    schedule=schedule)
    oacc, oout, oids, ocnt = cpu_render(oracle, 256, 256, True, 3, scene=rubik, cam=inside, max_depth=8)
    assert np.array_equal(ids, oids)
    assert_bit_exact(acc, oacc)
    assert np.array_equal(out, oout)
    assert cnt["segments"] == ocnt["segments"] and cnt["shadow_rays"] == ocnt["shadow_rays"]
