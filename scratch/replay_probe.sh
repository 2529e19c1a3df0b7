#!/bin/bash
# builds scratch/liboracle_probe.so: the oracle with Intersects() replaced by the candidates+replay experiment
set -e
cd "$(dirname "$0")"
sed 's/^uint32_t Intersects(const Inv& g, uint32_t bvh_start_index/uint32_t IntersectsOrig(const Inv\& g, uint32_t bvh_start_index/' ../oracle/shader_oracle.cpp > _probe_src.cpp
python - <<'P'
s = open('_probe_src.cpp').read()
# forward declaration right after the Inv struct's helper, and the probe body before the first use site (brdf section)
marker = '// ---- brdf.glsl'
i = s.index(marker)
decl = 'uint32_t Intersects(const Inv& g, uint32_t bvh_start_index, V3 ray_origin, V3 ray_dir, float& intersection_distance, V3& tri_norm);\n'
s = s[:i] + decl + s[i:]
# close the anonymous namespace, include the probe (which reopens it at its end)
j = s.index('uint32_t IntersectsOrig(')
k = s.index(decl)
s = s[:k] + decl + '#include "replay_probe.inc"\n' + s[k + len(decl):]
s = '#include <algorithm>\n#include <map>\n#include <mutex>\n#include <stdio.h>\n' + s
open('_probe_src.cpp', 'w').write(s)
P
g++ -O2 -std=c++17 -ffp-contract=off -mavx2 -mfma -pthread -fPIC -w -I../include -I../oracle -shared -o liboracle_probe.so _probe_src.cpp
echo built
