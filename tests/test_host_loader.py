"""CPU tests of the host C++ side (simple-ray-tracer_b200/host): OBJ/MTL loader, BVH builder, scene flattening.

The bar is BYTE equality with the reference's own loader + bvh.h: tests/golden/rubik_scene.npz was written by
oracle/_ref/ref_scene_dump (the reference's sources compiled where they lie); where /root/reference is present the
same binary is also run on generated OBJ files so the builder is compared on inputs the reference never shipped
(quads, one-sided splits, duplicate centroids, faces before the first usemtl, repeated newmtl names).
"""
import os
import struct
import subprocess

import numpy as np
import pytest

import scenes

HOST = os.path.join(scenes.ROOT, "simple-ray-tracer_b200", "host")
DUMP = os.path.join(HOST, "srt_scene_dump")
REF_DUMP = os.path.join(scenes.ROOT, "oracle", "_ref", "ref_scene_dump")
REF_ROOT = "/root/reference"
NAMES = ("bvhs", "nodes", "materials", "triangles", "vertices")


@pytest.fixture(scope="module")
def host_dump():
    if not os.path.exists(DUMP):
        subprocess.run(["make", "-s", "-C", HOST, "srt_scene_dump"], check=True)
    return DUMP


def read_dump(path):
    b = open(path, "rb").read()
    assert b[:8] == b"SRTREF01"
    off, out = 8, []
    for _ in range(5):
        (n,) = struct.unpack_from("<Q", b, off)
        off += 8
        out.append(b[off:off + n])
        off += n
    return dict(zip(NAMES, out))


def canonical(bufs):
    """Zeroes the bytes the reference leaves uninitialised (struct padding, handle of untextured materials)."""
    import make_scene_golden as M  # tests/golden
    bvhs = np.frombuffer(bufs["bvhs"], M.BVH_DT).copy()
    mats = np.frombuffer(bufs["materials"], M.MAT_DT).copy()
    verts = np.frombuffer(bufs["vertices"], M.VERT_DT).copy()
    bvhs["_pad"] = 0
    mats["_pad"] = 0
    mats["handle"][mats["use_texture"] == 0] = 0
    verts["_pad0"] = 0
    verts["_pad1"] = 0
    return {"bvhs": bvhs.tobytes(), "nodes": bufs["nodes"], "materials": mats.tobytes(),
            "triangles": bufs["triangles"], "vertices": verts.tobytes()}


def run_host(dump, objects_dir, name, tmp_path, expect_ok=True):
    out = str(tmp_path / (name + "_host.bin"))
    r = subprocess.run([dump, str(objects_dir), name, out], capture_output=True, text=True)
    if expect_ok:
        assert r.returncode == 0, r.stderr
        return read_dump(out)
    return r


def run_reference(root, name, tmp_path):
    out = str(tmp_path / (name + "_ref.bin"))
    subprocess.run([REF_DUMP, str(root), name, out], check=True, capture_output=True)
    return read_dump(out)


needs_reference = pytest.mark.skipif(not (os.path.isdir(REF_ROOT) and os.path.exists(REF_DUMP)),
                                     reason="reference mount / oracle/_ref not present on this box")


@needs_reference
def test_rubik_arrays_equal_reference_golden_bytes(host_dump, tmp_path):
    import sys
    sys.path.insert(0, scenes.GOLDEN)
    got = canonical(run_host(host_dump, os.path.join(REF_ROOT, "objects"), "Rubik", tmp_path))
    want = scenes.rubik_scene()
    for k in NAMES:
        assert got[k] == want[k].tobytes(), k
    nodes = want["nodes"]
    assert len(nodes) == 427 and int((nodes["count"] > 0).sum()) == 214 and int(nodes["count"].max()) == 13


def write_obj(folder, name, rng, n_tri, n_quad, cluster=False, pre_usemtl_faces=0, dup_newmtl=False):
    d = folder / "objects" / name
    d.mkdir(parents=True)
    mats = ["red", "green", "blue", "grey"]
    with open(d / "lib.mtl", "w") as f:
        f.write("# generated\n")
        for i, m in enumerate(mats):
            f.write("newmtl %s\nKd %.4f %.4f %.4f\nKs 0.1 0.2 %.3f\nNs %d\nillum 2\n\n" % (m, 0.2 * i, 0.9 - 0.1 * i, 0.3 + 0.1 * i, 0.05 * i, 5 + 7 * i))
        if dup_newmtl:
            f.write("newmtl red\nKd 0.5 0.5 0.5\nNs 77\n")  # Q24: overwrites the material that was current (grey)
    lines = ["mtllib lib.mtl", "o thing"]
    verts = []

    def v(p):
        verts.append(p)
        return len(verts)

    faces = []
    for i in range(n_tri + n_quad):
        if cluster and i % 3 == 0:
            c = np.array([1.0, 2.0, 3.0])  # many identical centroids: one-sided midpoint splits, big leaves
            e = rng.uniform(-0.25, 0.25, (2, 3))
            pts = [c - e[0] - e[1], c + 2 * e[0] - e[1], c - e[0] + 2 * e[1]]
            if i >= n_tri:
                pts.append(c + e[0] + e[1])
        else:
            c = rng.uniform(-20, 20, 3)
            pts = [c + rng.uniform(-1, 1, 3) for _ in range(4 if i >= n_tri else 3)]
        faces.append([v(p) for p in pts])
    for p in verts:
        lines.append("v %.6f %.6f %.6f" % tuple(p))
    lines.append("vt 0.5 0.5")
    lines.append("vn 0 1 0")
    order = rng.permutation(len(faces))
    for k, fi in enumerate(order):
        if k == pre_usemtl_faces or (k > pre_usemtl_faces and k % 97 == 0):
            lines.append("usemtl %s" % mats[(k // 97) % len(mats)])
        f = faces[fi]
        style = k % 3
        lines.append("f " + " ".join(("%d" % a) if style == 0 else ("%d/1/1" % a) if style == 1 else ("%d//1" % a) for a in f))
    with open(d / (name + ".obj"), "w") as fobj:
        fobj.write("\r\n".join(lines) + "\r\n")  # CRLF like the reference's Rubik.obj
    return d


@needs_reference
@pytest.mark.parametrize("case", ["soup", "clustered", "quirks"])
def test_generated_objs_equal_reference_loader_bytes(host_dump, tmp_path, case):
    rng = np.random.default_rng({"soup": 1, "clustered": 2, "quirks": 3}[case])
    if case == "soup":
        write_obj(tmp_path, "gen", rng, 3000, 500)
    elif case == "clustered":
        write_obj(tmp_path, "gen", rng, 900, 300, cluster=True)
    else:
        write_obj(tmp_path, "gen", rng, 200, 40, pre_usemtl_faces=17, dup_newmtl=True)
    got = canonical(run_host(host_dump, tmp_path / "objects", "gen", tmp_path))
    want = canonical(run_reference(tmp_path, "gen", tmp_path))
    for k in NAMES:
        assert got[k] == want[k], k
    import sys
    sys.path.insert(0, scenes.GOLDEN)
    import make_scene_golden as M
    nodes = np.frombuffer(got["nodes"], M.NODE_DT)
    assert len(nodes) > 100
    if case == "clustered":
        assert nodes["count"].max() > 2  # one-sided splits leave leaves with more than two triangles (Q25)


def test_loader_without_reference_properties(host_dump, tmp_path):
    """Runs everywhere: structure of the arrays for a generated OBJ."""
    import sys
    sys.path.insert(0, scenes.GOLDEN)
    import make_scene_golden as M
    rng = np.random.default_rng(5)
    write_obj(tmp_path, "gen", rng, 500, 100)
    bufs = run_host(host_dump, tmp_path / "objects", "gen", tmp_path)
    nodes = np.frombuffer(bufs["nodes"], M.NODE_DT)
    tris = np.frombuffer(bufs["triangles"], M.TRI_DT)
    verts = np.frombuffer(bufs["vertices"], M.VERT_DT)
    bvhs = np.frombuffer(bufs["bvhs"], M.BVH_DT)
    assert len(tris) == 500 + 2 * 100 and len(verts) == 3 * len(tris)  # quads fan into two, corners never shared
    assert len(bvhs) == 1 and bvhs[0]["count"] == len(nodes) and np.array_equal(bvhs[0]["frame"], np.eye(4, dtype=np.float32).ravel())
    assert (verts["texture"] == 0).all()  # Q6
    leaves = nodes[nodes["count"] > 0]
    assert leaves["count"].sum() == len(tris)  # every triangle in exactly one leaf
    covered = np.zeros(len(tris), int)
    for lf in leaves:
        covered[lf["index"]:lf["index"] + lf["count"]] += 1
    assert (covered == 1).all()
    inner = nodes[nodes["count"] == 0]
    assert (inner["index"] + 1 < len(nodes)).all() and len(nodes) == 2 * len(leaves) - 1
    # every leaf box bounds its triangles
    P = verts["vertex"][tris["v"]]  # [T,3,3]
    for lf in leaves[:50]:
        pts = P[lf["index"]:lf["index"] + lf["count"]].reshape(-1, 3)
        assert (pts >= lf["min"] - 0).all() and (pts <= lf["max"] + 0).all()


def test_missing_obj_reports_error_getting_geo(host_dump, tmp_path):
    (tmp_path / "objects" / "nothing").mkdir(parents=True)
    r = run_host(host_dump, tmp_path / "objects", "nothing", tmp_path, expect_ok=False)
    assert r.returncode == 1 and "error getting geo" in r.stderr  # model_loader.cpp:28-29


def test_two_models_flatten_with_rebased_indices(host_dump, tmp_path):
    """UploadModelDataToGPU over several models (gpu_loader.cpp:63-133): arrays appended in order, triangle corners
    rebased by the vertices before, materials by the materials before, leaves by the triangles before, inner nodes by
    the nodes before; one BVH range per model."""
    import sys
    sys.path.insert(0, scenes.GOLDEN)
    import make_scene_golden as M
    write_obj(tmp_path, "first", np.random.default_rng(11), 300, 60)
    write_obj(tmp_path, "second", np.random.default_rng(12), 120, 10, cluster=True)
    one = {n: run_host(host_dump, tmp_path / "objects", n, tmp_path) for n in ("first", "second")}
    both = run_host(host_dump, tmp_path / "objects", "first,second", tmp_path)
    dts = {"bvhs": M.BVH_DT, "nodes": M.NODE_DT, "materials": M.MAT_DT, "triangles": M.TRI_DT, "vertices": M.VERT_DT}
    a = {k: np.frombuffer(one["first"][k], dts[k]) for k in NAMES}
    b = {k: np.frombuffer(one["second"][k], dts[k]).copy() for k in NAMES}
    got = {k: np.frombuffer(both[k], dts[k]) for k in NAMES}
    b["triangles"]["v"] += len(a["vertices"])
    b["triangles"]["material"] += len(a["materials"])
    leaf = b["nodes"]["count"] > 0
    b["nodes"]["index"][leaf] += len(a["triangles"])
    b["nodes"]["index"][~leaf] += len(a["nodes"])
    b["bvhs"]["first_index"] += len(a["nodes"])
    for k in NAMES:
        assert np.concatenate([a[k], b[k]]).tobytes() == got[k].tobytes(), k
    assert len(got["bvhs"]) == 2 and got["bvhs"]["count"].tolist() == [len(a["nodes"]), len(b["nodes"])]


def decode_png_with_host(path, tmp_path):
    exe = os.path.join(HOST, "srt_png_dump")
    if not os.path.exists(exe):
        subprocess.run(["make", "-s", "-C", HOST, "srt_png_dump"], check=True)
    out = str(tmp_path / "png.raw")
    r = subprocess.run([exe, str(path), out], capture_output=True, text=True)
    if r.returncode != 0:
        return None, r.stderr
    b = open(out, "rb").read()
    w, h = struct.unpack_from("<II", b, 0)
    return np.frombuffer(b, np.uint8, w * h * 4, 8).reshape(h, w, 4), ""


@pytest.mark.parametrize("channels", [1, 2, 3, 4, "palette"])
def test_png_decoder_every_colour_type_and_filter(tmp_path, channels):
    """The texture option's decoder (asset_utils/png_decode.h) against files written from known pixels with every
    scanline filter, two IDAT chunks."""
    rng = np.random.default_rng(3)
    H, W = 11, 7
    path = tmp_path / "t.png"
    if channels == "palette":
        pal = rng.integers(0, 256, (5, 3), dtype=np.uint8)
        idx = rng.integers(0, 5, (H, W), dtype=np.uint8)
        scenes.write_png(path, idx, colour_type=3, filters=[0, 1, 2, 3, 4], palette=pal)
        want = np.concatenate([pal[idx], np.full((H, W, 1), 255, np.uint8)], axis=2)
    else:
        img = rng.integers(0, 256, (H, W, channels), dtype=np.uint8)
        scenes.write_png(path, img, filters=[4, 3, 2, 1, 0])
        if channels == 1:
            want = np.concatenate([np.repeat(img, 3, 2), np.full((H, W, 1), 255, np.uint8)], 2)
        elif channels == 2:
            want = np.concatenate([np.repeat(img[:, :, :1], 3, 2), img[:, :, 1:]], 2)
        elif channels == 3:
            want = np.concatenate([img, np.full((H, W, 1), 255, np.uint8)], 2)
        else:
            want = img
    got, err = decode_png_with_host(path, tmp_path)
    assert got is not None, err
    assert np.array_equal(got, want)


def test_png_decoder_refuses_what_it_cannot_read(tmp_path):
    (tmp_path / "no.png").write_bytes(b"not a png at all, not even close........")
    got, err = decode_png_with_host(tmp_path / "no.png", tmp_path)
    assert got is None and "not a PNG" in err
    got, err = decode_png_with_host(tmp_path / "missing.png", tmp_path)
    assert got is None and "cannot open" in err
