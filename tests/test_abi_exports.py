"""CPU tests of the drop-in boundary: libsrt_b200.so loads without a GPU and exports every entry point that
include/srt_b200.h declares; the Python view's struct mirrors have the sizes of the C structs; and without a CUDA
device the library refuses to create a context (there is no CPU fallback to fall into)."""
import ctypes as C
import os
import re
import subprocess

import pytest

import scenes

HEADER = os.path.join(scenes.ROOT, "include", "srt_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(srt_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_boundary():
    names = declared_functions()
    for must in ("srt_create", "srt_destroy", "srt_upload_scene", "srt_update_frame", "srt_upload_noise", "srt_set_lights",
                 "srt_set_params", "srt_resize", "srt_render", "srt_read_output", "srt_read_accum", "srt_trace_rays"):
        assert must in names


def test_library_exports_every_declared_symbol(pkg):
    lib = C.CDLL(pkg.LIB_PATH)
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, missing
    assert sorted(pkg.EXPORTS) == declared_functions()
    lib.srt_version.restype = C.c_char_p
    assert b"sm_100a" in lib.srt_version()


def test_option_numbers_match_the_header(pkg):
    """Every SRT_OPT_* of the header has the same number in the Python view, and no two options share one."""
    src = open(HEADER).read()
    opts = {k: int(v) for k, v in re.findall(r"#define\s+SRT_OPT_([A-Z0-9_]+)\s+(\d+)", src)}
    assert len(opts) >= 17 and len(set(opts.values())) == len(opts), opts
    for name, number in opts.items():
        assert getattr(pkg, "OPT_" + name) == number, name


def test_struct_mirrors_match_the_c_layouts(pkg, tmp_path):
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "srt_b200.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                   "sizeof(srt_params),sizeof(srt_counters),sizeof(srt_bvh),sizeof(srt_bvh_node),sizeof(srt_material),"
                   "sizeof(srt_triangle),sizeof(srt_vertex),sizeof(srt_light),sizeof(srt_ray));return 0;}\n")
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-std=c99", "-I", os.path.join(scenes.ROOT, "include"), str(src), "-o", str(exe)], check=True)
    sizes = [int(v) for v in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert sizes == [C.sizeof(pkg.Params), C.sizeof(pkg.Counters), pkg.BVH_DT.itemsize, pkg.NODE_DT.itemsize,
                     pkg.MAT_DT.itemsize, pkg.TRI_DT.itemsize, pkg.VERT_DT.itemsize, pkg.LIGHT_DT.itemsize,
                     pkg.RAY_DT.itemsize]
    assert sizes[2:] == [80, 32, 48, 16, 32, 32, 32]  # std430 sizes of raytrace_types.glsl


def test_no_gpu_means_no_context(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is visible here")
    with pytest.raises(pkg.SrtError) as e:
        pkg.Context(0)
    assert "no CPU fallback" in str(e.value)


def test_product_does_not_reach_into_the_oracle():
    """Nothing under simple-ray-tracer_b200/ may include, link or import oracle/ (the oracle is the checker)."""
    root = os.path.join(scenes.ROOT, "simple-ray-tracer_b200")
    bad = []
    for d, _, files in os.walk(root):
        for f in files:
            if f.endswith((".cu", ".cuh", ".h", ".hpp", ".cpp", ".py", "Makefile")):
                text = open(os.path.join(d, f), errors="ignore").read()
                if re.search(r"oracle[/_.]|liboracle|shader_oracle", text):
                    bad.append(os.path.join(d, f))
    assert not bad, bad
