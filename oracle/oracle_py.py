"""ctypes binding of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

BVH_DT = np.dtype([("first_index", "<u4"), ("count", "<u4"), ("_pad", "<u4", 2), ("frame", "<f4", 16)])
NODE_DT = np.dtype([("min", "<f4", 3), ("index", "<u4"), ("max", "<f4", 3), ("count", "<u4")])
MAT_DT = np.dtype([("diffuse", "<f4", 3), ("specular_ex", "<f4"), ("specular", "<f4", 3), ("use_texture", "<u4"),
                   ("handle", "<u4", 2), ("_pad", "<u4", 2)])
TRI_DT = np.dtype([("v", "<u4", 3), ("material", "<u4")])
VERT_DT = np.dtype([("vertex", "<f4", 3), ("_pad0", "<f4"), ("texture", "<f4", 2), ("_pad1", "<f4", 2)])
LIGHT_DT = np.dtype([("position", "<f4", 3), ("intensity", "<f4"), ("color", "<f4", 3), ("pad1", "<f4")])
RAY_DT = np.dtype([("origin", "<f4", 3), ("_pad0", "<f4"), ("direction", "<f4", 3), ("dist", "<f4")])


class Params(C.Structure):
    _fields_ = [("show_model", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
                ("accum_frames", C.c_int32), ("reset_accum", C.c_int32),
                ("camera_origin", C.c_float * 3), ("camera_direction", C.c_float * 3),
                ("camera_up", C.c_float * 3), ("camera_right", C.c_float * 3),
                ("light_count", C.c_int32), ("bvh_count", C.c_uint32), ("max_depth", C.c_int32)]


class Counters(C.Structure):
    _fields_ = [("segments", C.c_uint64), ("shadow_rays", C.c_uint64), ("nodes", C.c_uint64), ("tris", C.c_uint64),
                ("accepted", C.c_uint64), ("noise_lookups", C.c_uint64), ("capped_paths", C.c_uint64),
                ("nan_pixels", C.c_uint64), ("path_len_hist", C.c_uint64 * 16)]

    def as_dict(self):
        d = {k: int(getattr(self, k)) for k, _ in self._fields_ if k != "path_len_hist"}
        d["path_len_hist"] = [int(v) for v in self.path_len_hist]
        return d


def build():
    subprocess.run(["make", "-s", "-C", HERE, "liboracle.so"], check=True)


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _lib = C.CDLL(path)
        _lib.oracle_render.restype = C.c_int
        _lib.oracle_trace_rays.restype = C.c_int
        for name in ("oracle_sin", "oracle_cos", "oracle_linear_to_srgb"):
            getattr(_lib, name).restype = C.c_float
            getattr(_lib, name).argtypes = [C.c_float]
        for name in ("oracle_pow", "oracle_rand_float"):
            getattr(_lib, name).restype = C.c_float
            getattr(_lib, name).argtypes = [C.c_float, C.c_float]
    return _lib


_ref_lib = None
SHADER_REF_PATH = os.path.join(HERE, "_ref", "libshader_ref.so")


def shader_ref_available():
    """True when oracle/_ref/libshader_ref.so (the reference's own shader text compiled as C++, oracle/shader_ref.cpp)
    has been built -- in the build container from /root/reference, on the GPU box from the prebuilt file."""
    return os.path.exists(SHADER_REF_PATH)


def ref_lib():
    global _ref_lib
    if _ref_lib is None:
        _ref_lib = C.CDLL(SHADER_REF_PATH)
        _ref_lib.shader_ref_render.restype = C.c_int
    return _ref_lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def make_params(width, height, cam, show_model, accum_frames=1, light_count=0, bvh_count=1, max_depth=0, reset=0):
    p = Params()
    p.show_model, p.width, p.height = int(show_model), width, height
    p.accum_frames, p.reset_accum = accum_frames, reset
    p.camera_origin[:] = cam["origin"]
    p.camera_direction[:] = cam["direction"]
    p.camera_up[:] = cam["up"]
    p.camera_right[:] = cam["right"]
    p.light_count, p.bvh_count, p.max_depth = light_count, bvh_count, max_depth
    return p


def render(params, n_frames, scene, noise_unit, noise_uniform, lights, accum=None, want_out=True, want_ids=True,
           rows=None, n_threads=0, impl="oracle"):
    """Returns (accum[H,W,4] f32, out[H,W,4] u8 or None, ids[H,W] u32 or None, counters dict).
    impl = "oracle": the hand-written restatement (shader_oracle.cpp); "shader_ref": the reference's shader text
    compiled as C++ (shader_ref.cpp) -- same signature, same decrees."""
    W, H = params.width, params.height
    if accum is None:
        accum = np.zeros((H, W, 4), np.float32)  # decree Q8
    out = np.zeros((H, W, 4), np.uint8) if want_out else None
    ids = np.full((H, W), 0xFFFFFFFF, np.uint32) if want_ids else None
    cnt = Counters()
    y0, y1 = rows if rows else (0, H)
    s = scene or {}
    lights = np.ascontiguousarray(lights)
    fn = lib().oracle_render if impl == "oracle" else ref_lib().shader_ref_render
    rc = fn(C.byref(params), C.c_int(n_frames), _p(s.get("bvhs")), _p(s.get("nodes")),
            _p(s.get("materials")), _p(s.get("triangles")), _p(s.get("vertices")),
            _p(noise_unit), _p(noise_uniform), _p(lights), C.c_int(len(lights)),
            _p(accum), _p(out), _p(ids), C.c_int(y0), C.c_int(y1), C.c_int(n_threads),
            C.byref(cnt))
    if rc != 0:
        raise RuntimeError("%s render failed: %d" % (impl, rc))
    return accum, out, ids, cnt.as_dict()


_texture_keepalive = []


def set_textures(textures):
    """Textured-albedo option: [(material index, uint8 array [H,W,4], row 0 first), ...]; [] restores decree Q7."""
    global _texture_keepalive
    imgs = [np.ascontiguousarray(img, np.uint8) for _, img in textures]
    n = len(imgs)
    mat = (C.c_uint32 * max(n, 1))(*[m for m, _ in textures])
    w = (C.c_uint32 * max(n, 1))(*[i.shape[1] for i in imgs])
    h = (C.c_uint32 * max(n, 1))(*[i.shape[0] for i in imgs])
    ptr = (C.c_void_p * max(n, 1))(*[i.ctypes.data for i in imgs])
    _texture_keepalive = imgs
    if lib().oracle_set_textures(mat, w, h, ptr, C.c_int(n)) != 0:
        raise RuntimeError("oracle_set_textures failed")


def tex_bilinear(image, s, t):
    image = np.ascontiguousarray(image, np.uint8)
    s = np.ascontiguousarray(s, np.float32)
    t = np.ascontiguousarray(t, np.float32)
    out = np.zeros((len(s), 3), np.float32)
    lib().oracle_tex_bilinear(_p(image), C.c_uint32(image.shape[1]), C.c_uint32(image.shape[0]), _p(s), _p(t),
                              C.c_uint32(len(s)), _p(out))
    return out


def tex_uv(v0, v1, v2, p, uv6):
    a = [np.ascontiguousarray(x, np.float32) for x in (v0, v1, v2, p, uv6)]
    out = np.zeros(2, np.float32)
    lib().oracle_tex_uv(*[_p(x) for x in a], _p(out))
    return out


def trace_rays(scene, rays, bvh_count=1):
    rays = np.ascontiguousarray(rays)
    n = len(rays)
    hit = np.zeros(n, np.uint32)
    t = np.zeros(n, np.float32)
    cnt = Counters()
    rc = lib().oracle_trace_rays(_p(scene["bvhs"]), C.c_uint32(bvh_count), _p(scene["nodes"]), _p(scene["triangles"]),
                                 _p(scene["vertices"]), _p(rays), C.c_uint32(n), _p(hit), _p(t), C.byref(cnt))
    if rc != 0:
        raise RuntimeError("oracle_trace_rays failed")
    return hit, t, cnt.as_dict()
