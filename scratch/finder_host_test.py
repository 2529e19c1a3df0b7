import os, sys, tempfile, subprocess, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import scenes
which = sys.argv[1] if len(sys.argv) > 1 else "rubik"
tmp = tempfile.mkdtemp()
if which == "rubik": scene = scenes.rubik_scene()
elif which == "airplane":
    objdir = os.path.join(tmp, "objects"); scenes.write_airplane_standin(objdir); scene = scenes.load_obj_scene(objdir, "airplane_standin", tmp)
else: scene, _ = scenes.soup_scene(int(sys.argv[2]) if len(sys.argv) > 2 else 100000, tmp)
print({k: (v.shape, v.dtype) for k, v in scene.items() if hasattr(v, "shape")})
for k, f in (("bvhs", "bvhs"), ("nodes", "nodes"), ("triangles", "tris"), ("vertices", "verts")):
    np.ascontiguousarray(scene[k]).tofile(os.path.join(tmp, f + ".bin"))
sys.exit(subprocess.call([os.path.join(ROOT, "scratch", "finder_host_test"), tmp]))
