"""Per-wave trace/logic times of one launch: python scratch/wave_log.py <spp> [stripe_world]   (SRT_WAVE_LOG=1)"""
import sys, time, os
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
os.environ["SRT_WAVE_LOG"] = "1"
import numpy as np, scenes, __graft_entry__ as g
pkg = g.load_package()
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 64
world = int(sys.argv[2]) if len(sys.argv) > 2 else 1
ctx = pkg.Context(0)
W, H = 1920, 1080
unit, uniform = scenes.noise_tables(W, H)
ctx.upload_scene(scenes.rubik_scene()); ctx.upload_noise(unit, uniform); ctx.set_lights(scenes.model_lights()); ctx.resize(W, H)
ctx.set_option(pkg.OPT_TIME_KERNELS, 1)
if world > 1:
    ctx.set_stripe(0, world)
ctx.set_params(pkg.make_params(W, H, scenes.reset_camera(True), True, accum_frames=1, bvh_count=1))
for rep in range(3):
    ctx.render(spp); ctx.kernel_times() if rep < 2 else None
t0 = time.perf_counter(); ctx.render(spp); dt = time.perf_counter() - t0
t, l, w = ctx.kernel_times()
print("render %.2f ms  trace %.2f  logic %.2f  waves %d" % (dt * 1e3, t, l, w))
