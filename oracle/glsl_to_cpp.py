#!/usr/bin/env python3
"""glsl_to_cpp.py -- mechanical text transform: the reference's compute shader -> a C++ class body.  TEST INFRASTRUCTURE.

    python oracle/glsl_to_cpp.py /root/reference/shaders/raytrace_compute.glsl oracle/_ref/shader_gen.inc

Reads the reference's GLSL WHERE IT LIES (never copied into the repo; the output goes to oracle/_ref/, which is
git-ignored) and rewrites only what C++ cannot parse.  oracle/shader_ref.cpp includes the result inside
`struct Invocation { ... }`, so uniforms and buffers become data members and every GLSL function a member function,
compiled against oracle/shim/glsl/glsl_shim.h.  This is NOT a GLSL compiler and brings up no graphics stack.

Rules (each is local and syntactic; none looks at what a function computes):
  R1  `#include <f>` / "f"       -> the file's text, resolved relative to the including file, depth <= 32
                                    (what src/graphics/Shader.cpp:39-72 does before glShaderSource)
  R2  comments                   -> removed (so that commented-out declarations are not transformed)
  R3  `#version`, `#extension`, `layout(local_size...) in;`  -> removed
  R4  `layout(..) buffer B { T name[]; };`  -> `ssbo<T> name;`
  R5  `layout(..) [writeonly] uniform T name;` and `uniform T name;`  -> `T name;`
  R6  `inout T x` / `out T x` in parameter lists  -> `T& x`
  R7  `T[N] x` in parameter lists  -> `T x[N]`  (arrays are only read by their callees)
  R8  floating literals without a suffix -> `f` suffix (a GLSL literal is a 32-bit float, a C++ one a double)
  R9  swizzles `.xy .yz .xyz`    -> accessor calls `.xy() .yz() .xyz()` (all are r-values in these files)
  R10 `cos(phi).x`               -> `cos(phi)` (component selection on a scalar)
  R11 `int(` conversions         -> `toInt(` (the decree's int(): srt_detmath.h f2i_)
  R12 decree Q9: `while (true)`  -> `while (q9_continue())`  (hard cap of SRT_MAX_PATH_ITERS iterations, counted)
  R13 `settings.maxDepth = 5;`   -> `settings.maxDepth = maxDepthParam;` (srt_params.max_depth, default 5; configs[4]
                                    needs depth 8 and the shader hard-codes 5 at raytrace_compute.glsl:370)
  R14 probes: the DEFINITIONS of CheckHit, CheckLightOccluded, SphereHit, Intersects, IntersectsBox and
      IntersectsTriangle get the suffix `__glsl`; the callers keep the plain names and reach them through
      forwarding wrappers in shader_ref.cpp that only COUNT calls and note which primitive won the first CheckHit
"""
import os
import re
import sys

PROBED = ("CheckHit", "CheckLightOccluded", "SphereHit", "Intersects", "IntersectsBox", "IntersectsTriangle")


def resolve_includes(path, depth=0):
    if depth > 32:
        raise RuntimeError("header inclusion depth limit reached")
    rel = os.path.dirname(path)
    out = []
    rex = re.compile(r'^[ ]*#[ ]*include[ ]+["<](.*)[">].*')
    with open(path, "r") as f:
        for line in f.read().splitlines():
            m = rex.search(line)
            if m:
                out.append(resolve_includes(os.path.join(rel, m.group(1)), depth + 1))
            else:
                out.append(line)
    return "\n".join(out)


def strip_comments(text):
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    return re.sub(r"//[^\n]*", "", text)


def transform(text):
    text = strip_comments(text)                                                               # R2
    text = re.sub(r"^[ \t]*#[ \t]*(version|extension)[^\n]*$", "", text, flags=re.M)          # R3
    text = re.sub(r"layout\s*\([^)]*\)\s*in\s*;", "", text)
    text = re.sub(r"layout\s*\([^)]*\)\s*buffer\s+\w+\s*\{\s*(\w+)\s+(\w+)\s*\[\s*\]\s*;\s*\}\s*;",
                  r"ssbo<\1> \2;", text)                                                      # R4
    text = re.sub(r"layout\s*\([^)]*\)\s*(?:writeonly\s+|readonly\s+)?uniform\s+", "", text)  # R5
    text = re.sub(r"^([ \t]*)uniform\s+", r"\1", text, flags=re.M)
    text = re.sub(r"\b(?:inout|out)\s+(\w+)\s+(\w+)", r"\1& \2", text)                        # R6
    text = re.sub(r"\b(\w+)\[(\w+)\]\s+(\w+)(?=\s*[,)])", r"\1 \3[\2]", text)                 # R7
    text = re.sub(r"(?<![\w.])((?:\d+\.\d*|\.\d+)(?:[eE][+-]?\d+)?|\d+[eE][+-]?\d+)(?![\w.])", r"\1f", text)  # R8
    text = re.sub(r"\b((?:cos|sin)\(\w+\))\.x\b", r"\1", text)                                # R10
    text = re.sub(r"\.(xyz|xy|yz)\b(?!\s*\()", r".\1()", text)                                # R9
    text = re.sub(r"(?<![\w.])int\(", "toInt(", text)                                         # R11
    text = re.sub(r"while\s*\(\s*true\s*\)", "while (q9_continue())", text)                   # R12
    text, n = re.subn(r"settings\.maxDepth\s*=\s*5\s*;", "settings.maxDepth = maxDepthParam;", text)  # R13
    if n != 1:
        raise RuntimeError("R13: expected exactly one `settings.maxDepth = 5;`, found %d" % n)
    for name in PROBED:                                                                       # R14
        text, n = re.subn(r"^([ \t]*\w+[ \t]+)%s(\s*\([^;{]*\)\s*\{)" % name, r"\1%s__glsl\2" % name, text, flags=re.M)
        if n != 1:
            raise RuntimeError("R14: expected exactly one definition of %s, found %d" % (name, n))
    return text


def main():
    src, dst = sys.argv[1], sys.argv[2]
    text = transform(resolve_includes(src))
    os.makedirs(os.path.dirname(os.path.abspath(dst)), exist_ok=True)
    with open(dst, "w") as f:
        f.write("// GENERATED by oracle/glsl_to_cpp.py from %s -- do not edit, do not commit\n" % src)
        f.write("#undef NAN\n#undef INF\n#undef M_PI\n")
        f.write(text)
        f.write("\n")


if __name__ == "__main__":
    main()
