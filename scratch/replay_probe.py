"""SCRATCH: renders frames through the oracle whose Intersects() also runs the candidates+replay experiment."""
import ctypes as C, os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import oracle_py as O
import scenes
O.lib()
probe = C.CDLL(os.path.join(ROOT, "scratch", "liboracle_probe.so"))
probe.oracle_render.restype = C.c_int
probe.probe_config.argtypes = [C.c_float, C.c_int]
pad = float(sys.argv[2]) if len(sys.argv) > 2 else 1e-4
leaf = int(sys.argv[3]) if len(sys.argv) > 3 else 4
probe.probe_config(pad, leaf)
O._lib = probe  # same exports
which = sys.argv[1] if len(sys.argv) > 1 else "rubik"
W, H, frames = 480, 270, 2
tmp = tempfile.mkdtemp()
cam = scenes.reset_camera(True)
depth = 0
if which == "rubik":
    scene = scenes.rubik_scene()
elif which == "airplane":
    objdir = os.path.join(tmp, "objects")
    scenes.write_airplane_standin(objdir)
    scene = scenes.load_obj_scene(objdir, "airplane_standin", tmp)
elif which == "soup":
    scene, info = scenes.soup_scene(200000, tmp)
    cam = scenes.reset_camera(True, origin=(0.0, 0.0, 120.0)); depth = 8
elif which == "inside":
    scene = scenes.rubik_scene(); cam = scenes.look_camera((0.3, 9.1, 12.0), -70.0, -12.0); depth = 8
unit, uniform = scenes.noise_tables(W, H)
lights = scenes.model_lights()
p = O.make_params(W, H, cam, True, 1, len(lights), 1, depth)
acc, out, ids, cnt = O.render(p, frames, scene, unit, uniform, lights)
print(which, "pad", pad, "leaf", leaf, cnt)
probe.probe_report()
