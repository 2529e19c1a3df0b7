"""Device BVH build vs the host builder on the C5 soup: n triangles, v0 in [-50,50]^3, edges in [-.5,.5]^3."""
import json, os, struct, subprocess, sys, tempfile, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np, scenes, __graft_entry__ as g
pkg = g.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
rng = np.random.default_rng(0xC0FFEE)
v0 = rng.uniform(-50, 50, (n, 3)).astype(np.float32)
v1 = v0 + rng.uniform(-0.5, 0.5, (n, 3)).astype(np.float32)
v2 = v0 + rng.uniform(-0.5, 0.5, (n, 3)).astype(np.float32)
centre = ((v0 + v1) + v2) / np.float32(3.0)
lo = np.minimum(np.minimum(v0, v1), v2); hi = np.maximum(np.maximum(v0, v1), v2)
ctx = pkg.Context(0)
ctx.build_bvh(centre[:1000], lo[:1000], hi[:1000])
res = []
for _ in range(3):
    t0 = time.perf_counter()
    nodes, order, depth, ms = ctx.build_bvh(centre, lo, hi)
    res.append((ms, (time.perf_counter() - t0) * 1e3))
out = {"n_prims": n, "n_nodes": len(nodes), "depth": depth, "device_ms": [round(r[0], 2) for r in res],
       "e2e_ms": [round(r[1], 2) for r in res]}
if "--cpu" in sys.argv:
    exe = "simple-ray-tracer_b200/host/srt_bvh_dump"
    d = tempfile.mkdtemp()
    with open(d + "/in", "wb") as f:
        f.write(struct.pack("<I", n)); f.write(centre.tobytes()); f.write(lo.tobytes()); f.write(hi.tobytes())
    t0 = time.perf_counter()
    subprocess.run([exe, d + "/in", d + "/out"], check=True)
    out["host_tool_s_incl_file_io"] = round(time.perf_counter() - t0, 2)
    b = open(d + "/out", "rb").read()
    n_nodes, hdepth = struct.unpack_from("<II", b, 0)
    same = b[8:8 + 32 * n_nodes] == nodes.tobytes() and np.array_equal(np.frombuffer(b, np.uint32, n, 8 + 32 * n_nodes), order)
    out["bytes_equal_host"] = bool(same)
print(json.dumps(out))
