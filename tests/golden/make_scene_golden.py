"""Generates tests/golden/rubik_scene.npz from the REFERENCE's own scene-building code.

Runs oracle/_ref/ref_scene_dump (the reference's model_loader.cpp + gpu_loader.cpp + bvh.h compiled where
they lie, see oracle/Makefile) on objects/Rubik and stores the five buffers the reference would upload at
bindings 5..9 (src/main.cpp:199-204, src/asset_utils/gpu_loader.cpp:135-182).  Bytes the reference leaves
uninitialised (GPUMaterial.handle/_pad when use_texture == 0, PackedVertexData padding, GPUBVH padding) are
zeroed so the fixture is deterministic.  Needs /root/reference; run from the repo root:

    make -C oracle && python tests/golden/make_scene_golden.py
"""
import os
import struct
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.environ.get("SRT_REFERENCE", "/root/reference")

BVH_DT = np.dtype([("first_index", "<u4"), ("count", "<u4"), ("_pad", "<u4", 2), ("frame", "<f4", 16)])
NODE_DT = np.dtype([("min", "<f4", 3), ("index", "<u4"), ("max", "<f4", 3), ("count", "<u4")])
MAT_DT = np.dtype([("diffuse", "<f4", 3), ("specular_ex", "<f4"), ("specular", "<f4", 3), ("use_texture", "<u4"),
                   ("handle", "<u4", 2), ("_pad", "<u4", 2)])
TRI_DT = np.dtype([("v", "<u4", 3), ("material", "<u4")])
VERT_DT = np.dtype([("vertex", "<f4", 3), ("_pad0", "<f4"), ("texture", "<f4", 2), ("_pad1", "<f4", 2)])


def dump(name):
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_scene_dump")
    with tempfile.TemporaryDirectory() as d:
        out = os.path.join(d, "scene.bin")
        subprocess.run([exe, REF, name, out], check=True, stdout=subprocess.DEVNULL)
        b = open(out, "rb").read()
    assert b[:8] == b"SRTREF01"
    off, bufs = 8, []
    for _ in range(5):
        (n,) = struct.unpack_from("<Q", b, off)
        off += 8
        bufs.append(b[off:off + n])
        off += n
    bvhs = np.frombuffer(bufs[0], dtype=BVH_DT).copy()
    nodes = np.frombuffer(bufs[1], dtype=NODE_DT).copy()
    mats = np.frombuffer(bufs[2], dtype=MAT_DT).copy()
    tris = np.frombuffer(bufs[3], dtype=TRI_DT).copy()
    verts = np.frombuffer(bufs[4], dtype=VERT_DT).copy()
    bvhs["_pad"] = 0
    mats["_pad"] = 0
    mats["handle"][mats["use_texture"] == 0] = 0
    verts["_pad0"] = 0
    verts["_pad1"] = 0
    return bvhs, nodes, mats, tris, verts


if __name__ == "__main__":
    bvhs, nodes, mats, tris, verts = dump("Rubik")
    path = os.path.join(ROOT, "tests", "golden", "rubik_scene.npz")
    np.savez_compressed(path, bvhs=bvhs, nodes=nodes, materials=mats, triangles=tris, vertices=verts)
    print(path, len(bvhs), len(nodes), len(mats), len(tris), len(verts), os.path.getsize(path), "bytes")
