import sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np, bench, __graft_entry__ as g
pkg = g.load_package()
name = sys.argv[1] if len(sys.argv) > 1 else "airplane-standin"
inp = bench.workload_inputs(name, 1000000)
ctx = pkg.Context(0)
W, H = inp["W"], inp["H"]
def T(label, fn):
    ctx.sync(); t = time.perf_counter(); fn(); ctx.sync(); print("%-14s %8.2f ms" % (label, (time.perf_counter() - t) * 1e3))
for rep in range(3):
    T("upload_scene", lambda: ctx.upload_scene(inp["scene"]))
    T("upload_noise", lambda: ctx.upload_noise(inp["unit"], inp["uniform"]))
    T("set_lights", lambda: ctx.set_lights(inp["lights"]))
    if rep == 0: ctx.resize(W, H)
    ctx.set_params(pkg.make_params(W, H, inp["cam"], True, accum_frames=1, bvh_count=1, reset=1)); T("clear", lambda: ctx.render(1))
    ctx.set_params(pkg.make_params(W, H, inp["cam"], True, accum_frames=1, bvh_count=1, max_depth=inp["max_depth"])); T("render", lambda: ctx.render(inp["spp"]))
    T("read_output", lambda: ctx.read_output())
