import sys
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np, scenes, __graft_entry__ as g
pkg = g.load_package()
sched = int(sys.argv[1]) if len(sys.argv) > 1 else 1
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 1
spp = int(sys.argv[3]) if len(sys.argv) > 3 else 16
ctx = pkg.Context(0)
W, H = 1920, 1080
unit, uniform = scenes.noise_tables(W, H)
ctx.upload_scene(scenes.rubik_scene()); ctx.upload_noise(unit, uniform); ctx.set_lights(scenes.model_lights()); ctx.resize(W, H)
ctx.set_option(pkg.OPT_TRAVERSAL, mode); ctx.set_option(pkg.OPT_SCHEDULE, sched)
ctx.set_params(pkg.make_params(W, H, scenes.reset_camera(True), True, accum_frames=1, bvh_count=1))
warm = int(sys.argv[4]) if len(sys.argv) > 4 else 2
if warm: ctx.render(warm)
ctx.render(spp)
print(ctx.counters())
