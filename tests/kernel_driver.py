"""Every kernel of the library on small frames, each case compared with the oracle (not a test file: tests/
test_checked_build_gpu.py runs it in a subprocess against the bounds-checked build, SRT_LIB=.../libsrt_b200_checked.so).

    python tests/kernel_driver.py            # prints one line per case, the check flags, and ALL OK / SOME CASES DIFFER

compute-sanitizer is closed on the pool this was developed on ("runs under it have left GPUs needing a reset"), so the
kernels' own bounds checks (csrc/srt_device.cuh SRT_CHECK, compiled in with -DSRT_CHECKED) stand in for memcheck: every
traversal-stack push, every queue / list / slot / sample / candidate index.  The driver is small enough to run under a
sanitizer where one is available.  Kernels reached: clear_accum, wf_init
(<cull> and plain), wf_find (<outside>, <primary> and plain), wf_replay, wf_trace (overflow walk, and the plain walk with the finder
off), wf_logic, wf_tail, wf_resolve, pathtrace (spheres, per-path and regeneration), pathtrace_persistent, tonemap,
trace_rays (allocating and static-scratch entry points), pack_noise (pinned tables), detmath, the BVH builder's kernels,
the probes.
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
FLAGS = [0]  # OR of every context's check_flags

import __graft_entry__ as g  # noqa: E402
import scenes  # noqa: E402

pkg = g.load_package()
O = g.load_oracle()
W, H, FRAMES = 128, 72, 2


def check(name, ctx, show_model, scene, cam=None, max_depth=0, frames=FRAMES, bvh_count=1):
    unit, uniform = scenes.noise_tables(W, H)
    lights = scenes.model_lights() if show_model else scenes.sphere_lights()
    cam = cam or scenes.reset_camera(show_model)
    if scene is not None:
        ctx.upload_scene(scene)
    ctx.upload_noise(unit, uniform)
    ctx.set_lights(lights)
    ctx.resize(W, H)
    ctx.enable_primary_ids(True)
    nb = bvh_count if show_model else 0
    ctx.set_params(pkg.make_params(W, H, cam, show_model, accum_frames=1, bvh_count=nb, reset=1))
    ctx.render(1)
    ctx.set_params(pkg.make_params(W, H, cam, show_model, accum_frames=1, bvh_count=nb, max_depth=max_depth))
    ctx.render(frames)
    acc, ids = ctx.read_accum(), ctx.read_primary_ids()
    p = O.make_params(W, H, cam, show_model, 1, len(lights), nb, max_depth)
    oacc, _, oids, _ = O.render(p, frames, scene, unit, uniform, lights)
    ok = np.array_equal(ids, oids) and np.array_equal(acc.view(np.uint32), oacc.view(np.uint32))
    flags = ctx.counters()["check_flags"]
    FLAGS[0] |= flags
    print("%-46s %s  check_flags 0x%x" % (name, "ok" if ok else "DIFFERS", flags), flush=True)
    return ok


def main():
    good = True
    rubik = scenes.rubik_scene()
    with tempfile.TemporaryDirectory() as d:
        soup, _ = scenes.soup_scene(3000, d)
    inside = scenes.look_camera((0.3, 9.1, 12.0), -70.0, -12.0)
    variants = [
        ("wavefront defaults (finder, cull, tail)", {}),
        ("wavefront, waves to the end through the finder", {pkg.OPT_TAIL_PATHS: 0, pkg.OPT_FINDER: 1}),
        ("wavefront, camera rays made by the find kernel", {pkg.OPT_FUSED_PRIMARY: 1, pkg.OPT_TAIL_PATHS: 0, pkg.OPT_FINDER: 1}),
        ("wavefront, same, tail at 1 Ki paths", {pkg.OPT_FUSED_PRIMARY: 1, pkg.OPT_TAIL_PATHS: 1024, pkg.OPT_FINDER: 1}),
        ("wavefront, plain walk, no cull, no tail", {pkg.OPT_TAIL_PATHS: 0, pkg.OPT_FINDER: 0, pkg.OPT_PRIMARY_CULL: 0}),
        ("wavefront, small pool (slots regenerate)", {pkg.OPT_WAVE_SLOTS: 2048, pkg.OPT_TAIL_PATHS: 256}),
        ("wavefront, ordered traversal, finder off", {pkg.OPT_FINDER: 0, pkg.OPT_TRAVERSAL: pkg.TRAV_ORDERED}),
        ("wavefront, queues sorted by origin cell", {pkg.OPT_SORT_RAYS: 1, pkg.OPT_TAIL_PATHS: 1024}),
        ("wavefront, sorted queues, plain walk, no tail", {pkg.OPT_SORT_RAYS: 1, pkg.OPT_FINDER: 0, pkg.OPT_TAIL_PATHS: 0}),
        ("persistent schedule", {pkg.OPT_SCHEDULE: pkg.SCHED_PERSISTENT}),
        ("regeneration schedule", {pkg.OPT_SCHEDULE: pkg.SCHED_REGEN}),
        ("per-path schedule", {pkg.OPT_SCHEDULE: pkg.SCHED_PER_PATH}),
    ]
    for name, opts in variants:
        ctx = pkg.Context(0)
        ctx.set_option(pkg.OPT_TRAVERSAL, pkg.TRAV_REFERENCE)
        for k, v in opts.items():
            ctx.set_option(k, v)
        good &= check("rubik: " + name, ctx, True, rubik)
        ctx.close()
    ctx = pkg.Context(0)
    ctx.set_option(pkg.OPT_TRAVERSAL, pkg.TRAV_REFERENCE)
    ctx.set_option(pkg.OPT_FINDER, 1)
    ctx.set_option(pkg.OPT_TAIL_PATHS, 512)
    good &= check("rubik from inside, depth 8 (overflow rays)", ctx, True, rubik, cam=inside, max_depth=8)
    good &= check("soup 3 k, depth 8, finder forced", ctx, True, soup, cam=scenes.reset_camera(True, origin=(0.0, 0.0, 120.0)), max_depth=8)
    good &= check("spheres", ctx, False, None)
    ctx.set_option(pkg.OPT_FUSED_PRIMARY, 1)
    ctx.set_option(pkg.OPT_TAIL_PATHS, 0)
    good &= check("rubik from inside, camera rays by the find kernel", ctx, True, rubik, cam=inside, max_depth=8)
    ctx.set_option(pkg.OPT_FUSED_PRIMARY, 0)
    # the ray entry points, the probes, the decree kernel, the device BVH builder
    ctx.upload_scene(rubik)
    rays = np.zeros(777, pkg.RAY_DT)
    rng = np.random.default_rng(3)
    rays["origin"] = rng.uniform(-30, 30, (777, 3)).astype(np.float32)
    rays["direction"] = (np.array([0.0, 9.0, 3.0], np.float32) - rays["origin"]) + rng.normal(size=(777, 3)).astype(np.float32)
    rays["dist"] = np.float32(1e30)
    a, b = ctx.trace_rays(rays), ctx.trace_rays_static(rays)
    oh, ot, _ = O.trace_rays(rubik, rays)
    ok = np.array_equal(a[0], oh) and np.array_equal(b[0], oh) and np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32))
    print("%-46s %s" % ("trace_rays / trace_rays_static", "ok" if ok else "DIFFERS"), flush=True)
    good &= ok
    x = rng.uniform(-100, 100, 4096).astype(np.float32)
    ctx.eval_detmath(0, x)
    ctx.measure_gather(64 << 10, 1)
    ctx.measure_issue(1)
    c = rng.uniform(-1, 1, (5000, 3)).astype(np.float32)
    nodes, order, depth, _ = ctx.build_bvh(c, c - 0.01, c + 0.01)
    print("%-46s %s" % ("device BVH build (5000 prims, %d nodes)" % len(nodes), "ok"), flush=True)
    ctx.tonemap(2)
    FLAGS[0] |= ctx.counters()["check_flags"]
    ctx.reset_counters()
    ctx.check_selftest()
    print("library: %s; check_flags over all cases 0x%x; self-test raises 0x%x" % (
        pkg.lib().srt_version().decode(), FLAGS[0], ctx.counters()["check_flags"]), flush=True)
    ctx.close()
    print("ALL OK" if good else "SOME CASES DIFFER", flush=True)
    return 0 if good else 1


if __name__ == "__main__":
    sys.exit(main())
