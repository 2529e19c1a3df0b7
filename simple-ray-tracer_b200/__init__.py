"""Python view of libsrt_b200.so (the C-ABI of include/srt_b200.h) through ctypes.

This is plumbing for tests/, bench.py and __graft_entry__.py: it only marshals numpy arrays into the C-ABI
calls a C++ host (simple-ray-tracer_b200/host) makes.  There is no Python or CPU rendering path here; if the
shared library is missing or no B200 is visible the calls fail loudly.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# SRT_LIB picks another build of the same C-ABI, e.g. the bounds-checked one (libsrt_b200_checked.so) in tests
LIB_PATH = os.environ.get("SRT_LIB") or os.path.join(HERE, "libsrt_b200.so")
CHECKED_LIB_PATH = os.path.join(HERE, "libsrt_b200_checked.so")

OPT_TRAVERSAL, OPT_COUNT_TRAVERSAL, OPT_L2_PERSIST, OPT_BLOCK_THREADS, OPT_SCHEDULE = 1, 2, 3, 4, 5
OPT_WAVE_SLOTS, OPT_SAMPLE_BUDGET_MB, OPT_REFILL, OPT_TIME_KERNELS, OPT_TEXTURES, OPT_TAIL_PATHS, OPT_FINDER = 6, 7, 8, 9, 10, 11, 12
OPT_PRIMARY_CULL = 13
OPT_SPLIT = 14
OPT_SORT_RAYS = 16
OPT_DEVICE_RELAYOUT = 15
OPT_DEVICE_FINDER_BUILD = 17
OPT_OVERLAP = 18
OPT_FUSED_PRIMARY = 19
SPLIT_BANDS, SPLIT_SAMPLES = 0, 1
GROUP_ID_BYTES = 128
TRACE_STATIC_MAX_RAYS = 65536
SCHED_PER_PATH, SCHED_REGEN, SCHED_PERSISTENT, SCHED_WAVEFRONT = 0, 1, 2, 3
TRAV_REFERENCE, TRAV_ORDERED = 0, 1
NO_HIT = 0xFFFFFFFF
MAX_LIGHTS = 10

BVH_DT = np.dtype([("first_index", "<u4"), ("count", "<u4"), ("_pad", "<u4", 2), ("frame", "<f4", 16)])
NODE_DT = np.dtype([("min", "<f4", 3), ("index", "<u4"), ("max", "<f4", 3), ("count", "<u4")])
MAT_DT = np.dtype([("diffuse", "<f4", 3), ("specular_ex", "<f4"), ("specular", "<f4", 3), ("use_texture", "<u4"),
                   ("handle", "<u4", 2), ("_pad", "<u4", 2)])
TRI_DT = np.dtype([("v", "<u4", 3), ("material", "<u4")])
VERT_DT = np.dtype([("vertex", "<f4", 3), ("_pad0", "<f4"), ("texture", "<f4", 2), ("_pad1", "<f4", 2)])
LIGHT_DT = np.dtype([("position", "<f4", 3), ("intensity", "<f4"), ("color", "<f4", 3), ("pad1", "<f4")])
RAY_DT = np.dtype([("origin", "<f4", 3), ("_pad0", "<f4"), ("direction", "<f4", 3), ("dist", "<f4")])

EXPORTS = ["srt_create", "srt_destroy", "srt_last_error", "srt_set_stream", "srt_set_option", "srt_upload_scene",
           "srt_update_frame", "srt_upload_noise", "srt_set_lights", "srt_set_params", "srt_resize", "srt_set_rows", "srt_set_stripe",
           "srt_render", "srt_render_async", "srt_sync", "srt_read_output", "srt_read_accum", "srt_write_accum",
           "srt_tonemap", "srt_accum_device_ptr", "srt_enable_primary_ids", "srt_read_primary_ids", "srt_trace_rays", "srt_trace_rays_static", "srt_build_bvh", "srt_set_texture",
           "srt_get_counters", "srt_get_upload_info", "srt_get_kernel_times", "srt_get_finder_info", "srt_reset_counters", "srt_eval_detmath", "srt_measure_gather", "srt_measure_issue", "srt_check_selftest", "srt_group_unique_id", "srt_group_join",
           "srt_create_multi", "srt_get_collective_ms", "srt_version"]


class Params(C.Structure):  # srt_params, include/srt_types.h
    _fields_ = [("show_model", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
                ("accum_frames", C.c_int32), ("reset_accum", C.c_int32),
                ("camera_origin", C.c_float * 3), ("camera_direction", C.c_float * 3),
                ("camera_up", C.c_float * 3), ("camera_right", C.c_float * 3),
                ("light_count", C.c_int32), ("bvh_count", C.c_uint32), ("max_depth", C.c_int32)]


class Counters(C.Structure):  # srt_counters, include/srt_b200.h
    _fields_ = [("segments", C.c_uint64), ("shadow_rays", C.c_uint64), ("paths", C.c_uint64),
                ("capped_paths", C.c_uint64), ("kernel_launches", C.c_uint64), ("nodes", C.c_uint64),
                ("tris", C.c_uint64), ("culled_primaries", C.c_uint64), ("check_flags", C.c_uint64)]


def build():
    """Compiles libsrt_b200.so for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    subprocess.run(["make", "-s", "-C", HERE, "libsrt_b200.so"], check=True)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libsrt_b200.so is not built (run __graft_entry__.build()); there is no fallback path")
        _lib = C.CDLL(LIB_PATH)
        _lib.srt_last_error.restype = C.c_char_p
        _lib.srt_last_error.argtypes = [C.c_void_p]
        _lib.srt_version.restype = C.c_char_p
        _lib.srt_destroy.restype = None
        _lib.srt_destroy.argtypes = [C.c_void_p]
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class SrtError(RuntimeError):
    pass


def group_unique_id():
    """The id rank 0 hands to every rank's Context.group_join (srt_group_unique_id)."""
    buf = (C.c_ubyte * GROUP_ID_BYTES)()
    if lib().srt_group_unique_id(buf, C.c_uint64(GROUP_ID_BYTES)) != 0:
        raise SrtError(lib().srt_last_error(None).decode())
    return bytes(buf)


class Context:
    """One srt_ctx: the stand-in for the reference's GL context + compute program + bound buffers."""

    def __init__(self, device=0, devices=None):
        """device: one GPU (srt_create).  devices=[...]: one process driving several GPUs through a head context
        (srt_create_multi)."""
        self._h = C.c_void_p()
        self._lib = lib()
        if devices is not None:
            arr = (C.c_int * len(devices))(*devices)
            rc = self._lib.srt_create_multi(C.byref(self._h), arr, C.c_int32(len(devices)))
        else:
            rc = self._lib.srt_create(C.byref(self._h), C.c_int(device))
        if rc != 0:
            raise SrtError(self._lib.srt_last_error(None).decode())
        self.width = self.height = 0

    def group_join(self, rank, world, group_id):
        """One process per GPU: joins the group whose id rank 0 got from group_unique_id() (collective call)."""
        buf = (C.c_ubyte * GROUP_ID_BYTES).from_buffer_copy(bytes(group_id))
        self._ck(self._lib.srt_group_join(self._h, C.c_int32(rank), C.c_int32(world), buf, C.c_uint64(GROUP_ID_BYTES)))

    def collective_ms(self):
        ms, n = C.c_double(), C.c_uint64()
        self._ck(self._lib.srt_get_collective_ms(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def close(self):
        if self._h:
            self._lib.srt_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise SrtError(self._lib.srt_last_error(self._h).decode())

    def set_option(self, key, value):
        self._ck(self._lib.srt_set_option(self._h, C.c_int(key), C.c_int(value)))

    def set_stream(self, cuda_stream_ptr):
        self._ck(self._lib.srt_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def upload_scene(self, scene):
        a = {k: np.ascontiguousarray(scene[k]) for k in ("bvhs", "nodes", "materials", "triangles", "vertices")}
        assert a["bvhs"].dtype.itemsize == 80 and a["nodes"].dtype.itemsize == 32 and a["materials"].dtype.itemsize == 48
        assert a["triangles"].dtype.itemsize == 16 and a["vertices"].dtype.itemsize == 32
        self._ck(self._lib.srt_upload_scene(self._h, _p(a["bvhs"]), C.c_uint32(len(a["bvhs"])), _p(a["nodes"]),
                                            C.c_uint32(len(a["nodes"])), _p(a["materials"]),
                                            C.c_uint32(len(a["materials"])), _p(a["triangles"]),
                                            C.c_uint32(len(a["triangles"])), _p(a["vertices"]),
                                            C.c_uint32(len(a["vertices"]))))

    def update_frame(self, index, mat4_column_major):
        m = np.ascontiguousarray(mat4_column_major, np.float32).reshape(16)
        self._ck(self._lib.srt_update_frame(self._h, C.c_uint32(index), _p(m)))

    def upload_noise(self, unit, uniform):
        unit = np.ascontiguousarray(unit, np.float32)
        uniform = np.ascontiguousarray(uniform, np.float32)
        assert unit.shape == uniform.shape and unit.shape[-1] == 3
        self._ck(self._lib.srt_upload_noise(self._h, _p(unit), _p(uniform), C.c_uint32(unit.size // 3)))

    def set_lights(self, lights):
        lights = np.ascontiguousarray(lights)
        assert lights.dtype.itemsize == 32
        self._ck(self._lib.srt_set_lights(self._h, _p(lights), C.c_int32(len(lights))))

    def set_params(self, params):
        self._ck(self._lib.srt_set_params(self._h, C.byref(params)))

    def resize(self, width, height):
        self._ck(self._lib.srt_resize(self._h, C.c_int32(width), C.c_int32(height)))
        self.width, self.height = width, height

    def set_rows(self, y0, y1):
        self._ck(self._lib.srt_set_rows(self._h, C.c_int32(y0), C.c_int32(y1)))

    def set_stripe(self, rank, world):
        self._ck(self._lib.srt_set_stripe(self._h, C.c_int32(rank), C.c_int32(world)))

    def render(self, n_frames=1):
        self._ck(self._lib.srt_render(self._h, C.c_int32(n_frames)))

    def render_async(self, n_frames=1):
        self._ck(self._lib.srt_render_async(self._h, C.c_int32(n_frames)))

    def sync(self):
        self._ck(self._lib.srt_sync(self._h))

    def read_output(self, out=None):
        if out is None:
            out = np.empty((self.height, self.width, 4), np.uint8)
        self._ck(self._lib.srt_read_output(self._h, _p(out)))
        return out

    def read_accum(self, out=None):
        if out is None:
            out = np.empty((self.height, self.width, 4), np.float32)
        self._ck(self._lib.srt_read_accum(self._h, _p(out)))
        return out

    def write_accum(self, accum):
        accum = np.ascontiguousarray(accum, np.float32)
        assert accum.size == self.width * self.height * 4
        self._ck(self._lib.srt_write_accum(self._h, _p(accum)))

    def tonemap(self, accum_frames):
        self._ck(self._lib.srt_tonemap(self._h, C.c_int32(accum_frames)))

    def accum_device_ptr(self):
        ptr, n = C.c_void_p(), C.c_uint64()
        self._ck(self._lib.srt_accum_device_ptr(self._h, C.byref(ptr), C.byref(n)))
        return ptr.value, n.value

    def enable_primary_ids(self, enable=True):
        self._ck(self._lib.srt_enable_primary_ids(self._h, C.c_int(int(enable))))

    def read_primary_ids(self):
        ids = np.empty((self.height, self.width), np.uint32)
        self._ck(self._lib.srt_read_primary_ids(self._h, _p(ids)))
        return ids

    def trace_rays(self, rays, bvh_count=1):
        rays = np.ascontiguousarray(rays)
        assert rays.dtype.itemsize == 32
        hit = np.empty(len(rays), np.uint32)
        t = np.empty(len(rays), np.float32)
        self._ck(self._lib.srt_trace_rays(self._h, _p(rays), C.c_uint32(len(rays)), C.c_uint32(bvh_count), _p(hit), _p(t)))
        return hit, t

    def trace_rays_static(self, rays, bvh_count=1):
        """srt_trace_rays_static: the same results through the library's static device scratch (no allocation; at most
        TRACE_STATIC_MAX_RAYS rays)."""
        rays = np.ascontiguousarray(rays)
        assert rays.dtype.itemsize == 32
        n = len(rays)
        hit = np.zeros(n, np.uint32)
        t = np.zeros(n, np.float32)
        self._ck(self._lib.srt_trace_rays_static(self._h, _p(rays), C.c_uint32(n), C.c_uint32(bvh_count), _p(hit), _p(t)))
        return hit, t

    def set_texture(self, material_index, image):
        """image: uint8 [H, W, 4], row 0 first (takes effect with OPT_TEXTURES = 1)."""
        image = np.ascontiguousarray(image, np.uint8)
        assert image.ndim == 3 and image.shape[2] == 4
        self._ck(self._lib.srt_set_texture(self._h, C.c_uint32(material_index), _p(image), C.c_uint32(image.shape[1]),
                                           C.c_uint32(image.shape[0])))

    def build_bvh(self, centre, lo, hi):
        """Device build of the reference's BVH (bvh.h:40-148) from per-primitive centres and boxes.
        Returns (nodes [NODE_DT, tree-relative indices], order [uint32], depth, device_ms)."""
        centre, lo, hi = (np.ascontiguousarray(a, np.float32).reshape(-1, 3) for a in (centre, lo, hi))
        n = len(centre)
        assert len(lo) == n and len(hi) == n
        nodes = np.zeros(max(2 * n - 1, 1), NODE_DT)
        order = np.zeros(max(n, 1), np.uint32)
        n_nodes, depth, ms = C.c_uint32(), C.c_uint32(), C.c_float()
        self._ck(self._lib.srt_build_bvh(self._h, _p(centre), _p(lo), _p(hi), C.c_uint32(n), _p(nodes), C.byref(n_nodes),
                                         _p(order), C.byref(depth), C.byref(ms)))
        return nodes[:n_nodes.value].copy(), order[:n], depth.value, ms.value

    def counters(self):
        c = Counters()
        self._ck(self._lib.srt_get_counters(self._h, C.byref(c)))
        return {k: int(getattr(c, k)) for k, _ in Counters._fields_}

    def kernel_times(self):
        """(trace_ms, logic_ms, waves) of the wavefront launches since the last call (OPT_TIME_KERNELS must be on)."""
        t, l, w = C.c_double(), C.c_double(), C.c_uint64()
        self._ck(self._lib.srt_get_kernel_times(self._h, C.byref(t), C.byref(l), C.byref(w)))
        return t.value, l.value, w.value

    def finder_info(self):
        """(active, cost_reference, cost_finder): see srt_get_finder_info."""
        a, r, f = C.c_int32(), C.c_double(), C.c_double()
        self._ck(self._lib.srt_get_finder_info(self._h, C.byref(a), C.byref(r), C.byref(f)))
        return bool(a.value), r.value, f.value

    def measure_gather(self, n_bytes, chase):
        """GB/s of random 64 B record gathers from an n_bytes array (see srt_measure_gather)."""
        g = C.c_double()
        self._ck(self._lib.srt_measure_gather(self._h, C.c_uint64(n_bytes), C.c_int32(1 if chase else 0), C.byref(g)))
        return g.value

    def check_selftest(self):
        self._ck(self._lib.srt_check_selftest(self._h))

    def upload_info(self):
        """(on_device, device_ms) of the last upload_scene (see srt_get_upload_info)."""
        on, ms = C.c_int32(), C.c_float()
        self._ck(self._lib.srt_get_upload_info(self._h, C.byref(on), C.byref(ms)))
        return bool(on.value), ms.value

    def measure_issue(self, mix):
        """Thread-instructions per second (in units of 1e9) of un-fused fp32 arithmetic (see srt_measure_issue)."""
        g = C.c_double()
        self._ck(self._lib.srt_measure_issue(self._h, C.c_int32(mix), C.byref(g)))
        return g.value

    def reset_counters(self):
        self._ck(self._lib.srt_reset_counters(self._h))

    def eval_detmath(self, which, x, y=None):
        x = np.ascontiguousarray(x, np.float32)
        y = np.ascontiguousarray(y, np.float32) if y is not None else None
        out = np.empty_like(x)
        self._ck(self._lib.srt_eval_detmath(self._h, C.c_int(which), _p(x), _p(y), _p(out), C.c_uint32(x.size)))
        return out


def make_params(width, height, cam, show_model, accum_frames=1, light_count=-1, bvh_count=1, max_depth=0, reset=0):
    p = Params()
    p.show_model, p.width, p.height = int(show_model), width, height
    p.accum_frames, p.reset_accum = accum_frames, reset
    p.camera_origin[:] = cam["origin"]
    p.camera_direction[:] = cam["direction"]
    p.camera_up[:] = cam["up"]
    p.camera_right[:] = cam["right"]
    p.light_count, p.bvh_count, p.max_depth = light_count, bvh_count, max_depth
    return p
