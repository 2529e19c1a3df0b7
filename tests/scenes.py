"""Deterministic inputs shared by the parity tests, smoke() and bench.py (no reference access at run time).

Everything here is an INPUT of both the oracle and the CUDA path: noise tables (src/main.cpp:269-279 fills
them with std::rand(); here a PCG64 stream so that they do not depend on the libc), the light lists of
src/main.cpp:582-595, the camera basis of RayTracer::Camera::Reset (src/raytracer/camera.cpp:187-212).
"""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

LIGHT_DT = np.dtype([("position", "<f4", 3), ("intensity", "<f4"), ("color", "<f4", 3), ("pad1", "<f4")])


def _lights(rows):
    a = np.zeros(len(rows), LIGHT_DT)
    for i, (pos, col, inten) in enumerate(rows):
        a[i]["position"], a[i]["color"], a[i]["intensity"] = pos, col, inten
    return a


def model_lights():  # src/main.cpp:584-589
    return _lights([((1.0, 10.0, 10.0), (1.0, 1.0, 1.0), 50.0), ((-5.0, 15.0, 10.0), (1.0, 0.2, 0.2), 15.0),
                    ((5.0, 15.0, 10.0), (0.2, 1.0, 0.2), 15.0), ((-5.0, 5.0, 10.0), (0.2, 0.2, 1.0), 15.0),
                    ((5.0, 5.0, 10.0), (1.0, 1.0, 0.1), 15.0), ((0.0, 21.0, 17.0), (1.0, 1.0, 1.0), 50.0)])


def sphere_lights():  # src/main.cpp:593-594
    return _lights([((1.0, 2.0, 0.0), (1.0, 1.0, 1.0), 10.0), ((-2.5, 2.0, 0.0), (1.0, 1.0, 1.0), 3.0)])


def reset_camera(show_model, origin=None):
    """Camera::Reset + UpdateCameraVectors with yaw=-90, pitch=0: cosf(radians(-90)) = -4.371139e-08."""
    c = np.float32(-4.371139e-08)
    if origin is None:
        origin = (0.0, 9.0, 40.0) if show_model else (0.0, 1.0, 4.0)
    return {"origin": tuple(float(v) for v in origin), "direction": (float(c), 0.0, -1.0), "up": (0.0, 1.0, 0.0),
            "right": (1.0, 0.0, float(c))}


def look_camera(origin, yaw_deg, pitch_deg):
    """Camera::UpdateCameraVectors (src/raytracer/camera.cpp:120-136) in fp32 for arbitrary yaw/pitch."""
    f32 = np.float32
    yaw, pitch = np.radians(f32(yaw_deg), dtype=f32), np.radians(f32(pitch_deg), dtype=f32)
    front = np.array([np.cos(yaw) * np.cos(pitch), np.sin(pitch), np.sin(yaw) * np.cos(pitch)], f32)
    front = front / np.sqrt((front * front).sum(dtype=f32))
    right = np.cross(front, np.array([0, 1, 0], f32)).astype(f32)
    right = right / np.sqrt((right * right).sum(dtype=f32))
    up = np.cross(right, front).astype(f32)
    up = up / np.sqrt((up * up).sum(dtype=f32))
    return {"origin": tuple(float(v) for v in origin), "direction": tuple(float(v) for v in front),
            "up": tuple(float(v) for v in up), "right": tuple(float(v) for v in right)}


def noise_tables(width, height, seed=0x5EED):
    """(unit[N,3], uniform[N,3]) fp32: unit vectors by rejection from [-1,1]^3 then uniforms in [0,1)
    (same construction as Common::randomUnitVector / randomVec3, include/common/utils.h:30-51)."""
    n = width * height
    rng = np.random.Generator(np.random.PCG64(seed))
    unit = np.empty((n, 3), np.float32)
    filled = 0
    while filled < n:
        p = rng.random((int((n - filled) * 2.2) + 16, 3), dtype=np.float32) * np.float32(2) - np.float32(1)
        l2 = (p * p).sum(axis=1, dtype=np.float32)
        ok = (l2 > 1e-30) & (l2 <= 1)
        p = p[ok] / np.sqrt(l2[ok])[:, None]
        k = min(len(p), n - filled)
        unit[filled:filled + k] = p[:k]
        filled += k
    uniform = rng.random((n, 3), dtype=np.float32)
    return np.ascontiguousarray(unit), np.ascontiguousarray(uniform)


def rubik_scene():
    """Flattened Rubik scene exactly as the reference's loader uploads it (tests/golden/make_scene_golden.py)."""
    z = np.load(os.path.join(GOLDEN, "rubik_scene.npz"))
    return {k: np.ascontiguousarray(z[k]) for k in ("bvhs", "nodes", "materials", "triangles", "vertices")}


# ---- scenes built at run time through the host loader (simple-ray-tracer_b200/host/srt_scene_dump) ---------------
HOST_DIR = os.path.join(ROOT, "simple-ray-tracer_b200", "host")
BVH_DT = np.dtype([("first_index", "<u4"), ("count", "<u4"), ("_pad", "<u4", 2), ("frame", "<f4", 16)])
NODE_DT = np.dtype([("min", "<f4", 3), ("index", "<u4"), ("max", "<f4", 3), ("count", "<u4")])
MAT_DT = np.dtype([("diffuse", "<f4", 3), ("specular_ex", "<f4"), ("specular", "<f4", 3), ("use_texture", "<u4"),
                   ("handle", "<u4", 2), ("_pad", "<u4", 2)])
TRI_DT = np.dtype([("v", "<u4", 3), ("material", "<u4")])
VERT_DT = np.dtype([("vertex", "<f4", 3), ("_pad0", "<f4"), ("texture", "<f4", 2), ("_pad1", "<f4", 2)])


def read_scene_dump(path):
    """Container written by srt_scene_dump / srt_soup_dump: magic + five length-prefixed std430 arrays."""
    import struct
    with open(path, "rb") as f:
        assert f.read(8) == b"SRTREF01"
        out = []
        for dt in (BVH_DT, NODE_DT, MAT_DT, TRI_DT, VERT_DT):
            (n,) = struct.unpack("<Q", f.read(8))
            out.append(np.frombuffer(f.read(n), dtype=dt).copy())
    return dict(zip(("bvhs", "nodes", "materials", "triangles", "vertices"), out))


def concat_scenes(a, b):
    """Two flattened scenes as one upload, indices of the second rebased the way UploadModelDataToGPU does
    (gpu_loader.cpp:120-130)."""
    b = {k: v.copy() for k, v in b.items()}
    b["triangles"]["v"] += len(a["vertices"])
    b["triangles"]["material"] += len(a["materials"])
    leaf = b["nodes"]["count"] > 0
    b["nodes"]["index"][leaf] += len(a["triangles"])
    b["nodes"]["index"][~leaf] += len(a["nodes"])
    b["bvhs"]["first_index"] += len(a["nodes"])
    return {k: np.ascontiguousarray(np.concatenate([a[k], b[k]])) for k in a}


def _host_tool(name):
    import subprocess
    exe = os.path.join(HOST_DIR, name)
    if not os.path.exists(exe):
        subprocess.run(["make", "-s", "-C", HOST_DIR, name], check=True)
    return exe


def load_obj_scene(objects_dir, name, workdir, textures=False):
    """OBJ -> flattened scene through the product's own loader + BVH builder.  textures=True switches the loader's
    texture option on (vt lines honoured, map_Kd PNGs decoded)."""
    import subprocess
    out = os.path.join(workdir, name.replace(",", "+") + ".scene.bin")
    subprocess.run([_host_tool("srt_scene_dump"), objects_dir, name, out] + (["textures"] if textures else []), check=True,
                   capture_output=True)
    return read_scene_dump(out)


def write_png(path, img, colour_type=None, filters=None, palette=None):
    """Minimal PNG writer for tests: img uint8 [H,W] (grey or palette indices), [H,W,2], [H,W,3] or [H,W,4]; every
    scanline uses the filter type filters[y % len(filters)] (0-4, PNG specification section 9)."""
    import struct
    import zlib
    img = np.ascontiguousarray(img, np.uint8)
    if img.ndim == 2:
        img = img[:, :, None]
    H, W, ch = img.shape
    if colour_type is None:
        colour_type = {1: 0, 2: 4, 3: 2, 4: 6}[ch]
    filters = filters or [0]

    def paeth(a, b, c):
        p = a + b - c
        pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
        return a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)

    raw = bytearray()
    prev = [0] * (W * ch)
    for y in range(H):
        cur = img[y].reshape(-1).tolist()
        ft = filters[y % len(filters)]
        raw.append(ft)
        for i, v in enumerate(cur):
            a = cur[i - ch] if i >= ch else 0
            b = prev[i]
            c = prev[i - ch] if i >= ch else 0
            pred = [0, a, b, (a + b) // 2, paeth(a, b, c)][ft]
            raw.append((v - pred) & 255)
        prev = cur

    def chunk(kind, data):
        return struct.pack(">I", len(data)) + kind + data + struct.pack(">I", zlib.crc32(kind + data) & 0xFFFFFFFF)

    z = zlib.compress(bytes(raw), 6)
    half = len(z) // 2  # two IDAT chunks: the decoder must join them
    body = chunk(b"IHDR", struct.pack(">IIBBBBB", W, H, 8, colour_type, 0, 0, 0))
    if palette is not None:
        body += chunk(b"PLTE", np.ascontiguousarray(palette, np.uint8).tobytes())
    body += chunk(b"IDAT", z[:half]) + chunk(b"IDAT", z[half:]) + chunk(b"IEND", b"")
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + body)


def soup_scene(n_triangles, workdir, seed=0xC0FFEE):
    """BASELINE.json configs[4]: random triangle soup built by the host BVH builder."""
    import json
    import subprocess
    out = os.path.join(workdir, "soup_%d.scene.bin" % n_triangles)
    r = subprocess.run([_host_tool("srt_soup_dump"), str(n_triangles), out, hex(seed)], check=True, capture_output=True,
                       text=True)
    info = json.loads(r.stdout.strip().splitlines()[-1])
    scene = read_scene_dump(out)
    os.remove(out)
    return scene, info


def write_airplane_standin(objects_dir, name="airplane_standin", detail=96):
    """STAND-IN for objects/11803_Airplane_v1_l1 (its .obj is absent from the reference mount, .MISSING_LARGE_BLOBS:1):
    a procedurally generated closed airliner-shaped mesh -- fuselage of revolution, two swept wings, two engines,
    tailplane and fin -- about 15 units long standing on y=0 like the reference's models, with the SIX materials of the
    reference's 11803_Airplane_v1_l1.mtl (Kd 1.0 / 0.7843, Ks 0.54 / 0.279, Ns 51 / 30; map_Kd lines dropped: decree Q7
    shades textured materials with Kd anyway).  Triangle count = f(detail); detail=96 -> about 45 k.  Never reported under
    the Airplane name."""
    d = os.path.join(objects_dir, name)
    os.makedirs(d, exist_ok=True)
    mats = [("11803_Airplane_body", 51.0, 1.0, 0.54), ("11803_Airplane_wing_R", 30.0, 1.0, 0.279),
            ("11803_Airplane_wing_details_R", 30.0, 1.0, 0.279), ("11803_Airplane_tail", 30.0, 1.0, 0.279),
            ("11803_Airplane_wing_details_L", 30.0, 1.0, 0.279), ("11803_Airplane_wing_L", 30.0, 0.7843, 0.279)]
    with open(os.path.join(d, name + ".mtl"), "w") as f:
        for m, ns, kd, ks in mats:
            f.write("newmtl %s\n\tNs %.4f\n\tKd %.4f %.4f %.4f\n\tKs %.4f %.4f %.4f\n\n" % (m, ns, kd, kd, kd, ks, ks, ks))
    verts, groups = [], []

    def loft(material, rings):
        """Tube through a list of rings (each [n,3]), quads between consecutive rings.  Ends are not capped with
        discs or fans (clusters of tiny or long thin triangles make the reference's midpoint-split builder produce
        leaves with hundreds of triangles); bodies are closed by shrinking their last rings instead."""
        base = len(verts) + 1
        n = len(rings[0])
        for r in rings:
            verts.extend(r.tolist())
        faces = []
        for i in range(len(rings) - 1):
            for j in range(n):
                a, b = base + i * n + j, base + i * n + (j + 1) % n
                faces.append((a, b, b + n, a + n))
        groups.append((material, faces))

    def closed(rings, steps=3):
        """Adds rings that shrink the two ends towards their centroids (rounded tips)."""
        def tip(ring, nxt):
            c = ring.mean(axis=0)
            out = []
            for k in range(1, steps + 1):
                sc = np.cos(0.5 * np.pi * k / (steps + 0.35))
                out.append(c + (ring - c) * sc + (ring.mean(axis=0) - nxt.mean(axis=0)) * 0.35 * np.sin(0.5 * np.pi * k / (steps + 0.35)))
            return out
        head = tip(rings[0], rings[1])[::-1]
        tail = tip(rings[-1], rings[-2])
        return head + list(rings) + tail

    t = np.linspace(0, 2 * np.pi, detail, endpoint=False)
    # fuselage: revolution around z, z in [-7.5, 7.5], centre height 4; rings equally spaced in the ellipse parameter
    us = np.linspace(0.04, np.pi - 0.04, detail + 1)
    loft(mats[0][0], [np.stack([1.1 * np.sin(u) ** 0.7 * np.cos(t), 4.0 + 1.1 * np.sin(u) ** 0.7 * np.sin(t),
                                np.full_like(t, -7.5 * np.cos(u))], 1) for u in us])

    def wing(material, side, root_z, span, sweep, chord0, chord1, thick, y0, dihedral, sections):
        rings = []
        for s in np.linspace(0, 1, sections):
            chord = chord0 + (chord1 - chord0) * s
            cx = side * (0.9 + span * s)
            cz = root_z + sweep * s
            cy = y0 + dihedral * s
            rings.append(np.stack([np.full_like(t, cx), cy + 0.5 * thick * chord * np.sin(t),
                                   cz + 0.5 * chord * np.cos(t)], 1))
        loft(material, closed(rings))

    sec = max(8, detail // 2)
    wing(mats[1][0], +1, -0.5, 7.0, 2.5, 3.2, 1.0, 0.12, 3.6, 0.8, sec)
    wing(mats[5][0], -1, -0.5, 7.0, 2.5, 3.2, 1.0, 0.12, 3.6, 0.8, sec)
    wing(mats[3][0], +1, 6.0, 2.6, 1.0, 1.6, 0.7, 0.10, 4.3, 0.2, sec // 2)
    wing(mats[3][0], -1, 6.0, 2.6, 1.0, 1.6, 0.7, 0.10, 4.3, 0.2, sec // 2)
    # fin: vertical loft
    rings = []
    for s in np.linspace(0, 1, sec // 2):
        chord = 2.2 - 1.3 * s
        rings.append(np.stack([0.06 * chord * np.sin(t), np.full_like(t, 4.9 + 2.8 * s), 5.8 + 1.4 * s + 0.5 * chord * np.cos(t)], 1))
    loft(mats[3][0], closed(rings))
    # engines: short closed pods under the wings
    for side, m in ((+1, mats[2][0]), (-1, mats[4][0])):
        ez = np.linspace(-1.6, 0.4, max(6, detail // 6))
        loft(m, closed([np.stack([side * 3.4 + 0.45 * np.cos(t), 3.0 + 0.45 * np.sin(t), np.full_like(t, z)], 1) for z in ez]))

    # pose: fuselage along x (side view from the reference's camera at (0,9,40)), rolled 35 degrees towards the camera so
    # the wings' upper surfaces are seen, scaled to a 33-unit fuselage centred at height 9 -- the model then covers
    # about a fifth of the picture, as the reference's Rubik model does
    P = np.asarray(verts, np.float64) - np.array([0.0, 4.0, 0.0])
    P = np.stack([P[:, 2], P[:, 1], -P[:, 0]], 1)
    phi = np.radians(35.0)
    P = np.stack([P[:, 0], np.cos(phi) * P[:, 1] - np.sin(phi) * P[:, 2], np.sin(phi) * P[:, 1] + np.cos(phi) * P[:, 2]], 1)
    verts = (P * 2.2 + np.array([0.0, 9.0, 0.0])).tolist()

    n_tri = 0
    with open(os.path.join(d, name + ".obj"), "w") as f:
        f.write("# procedurally generated stand-in, see tests/scenes.py:write_airplane_standin\nmtllib %s.mtl\n" % name)
        for v in verts:
            f.write("v %.6f %.6f %.6f\n" % tuple(v))
        for m, faces in groups:
            f.write("usemtl %s\n" % m)
            for fc in faces:
                f.write("f " + " ".join(str(i) for i in fc) + "\n")
                n_tri += len(fc) - 2
    return name, n_tri
