import sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np, scenes, __graft_entry__ as g
pkg = g.load_package()
ctx = pkg.Context(0)
W, H = 1920, 1080
unit, uniform = scenes.noise_tables(W, H)
scene = scenes.rubik_scene()
ctx.upload_scene(scene); ctx.upload_noise(unit, uniform); ctx.set_lights(scenes.model_lights()); ctx.resize(W, H)
for mode in (0, 1):
    for sched, block in ((1, 64), (2, 64), (2, 128)):
        ctx.set_option(pkg.OPT_TRAVERSAL, mode); ctx.set_option(pkg.OPT_BLOCK_THREADS, block); ctx.set_option(pkg.OPT_SCHEDULE, sched)
        ctx.set_params(pkg.make_params(W, H, scenes.reset_camera(True), True, accum_frames=1, bvh_count=1))
        ctx.render(2)
        ctx.reset_counters()
        t = time.time(); ctx.render(16); dt = time.time() - t
        c = ctx.counters()
        print("mode", mode, "sched", sched, "block", block, "16spp ms", dt * 1e3, "Mseg/s", c["segments"] / dt / 1e6, "Mrays/s", (c["segments"] + c["shadow_rays"]) / dt / 1e6, c)
