// scratch: how often does a near-first traversal give a different closest hit than the shader's order?
#include "../oracle/shader_oracle.cpp"
#include <cstdio>
#include <random>
static float box2(V3 o, V3 d, const srt_bvh_node& n, bool& inside) {
  V3 inv = v3(1.0f) / d;
  V3 t0 = (ld3(n.min_bounds) - o) * inv, t1 = (ld3(n.max_bounds) - o) * inv;
  V3 tmn = min_(t0, t1), tmx = max_(t0, t1);
  float tn = max_(max_(tmn.x, tmn.y), tmn.z), tf = min_(min_(tmx.x, tmx.y), tmx.z);
  inside = (tn <= tf) && !(tn >= 0.0f);
  return tn <= tf ? (tn >= 0.0f ? tn : tf) : inf_();
}
static uint32_t ordered(const Inv& g, uint32_t root, V3 o, V3 d, float& best) {
  struct E { uint32_t n; float dist; } st[128]; int sp = 0;
  uint32_t hit = (uint32_t)-1;
  float dr = IntersectsBox(o, d, ld3(g.nodes[root].min_bounds), ld3(g.nodes[root].max_bounds));
  if (!(dr < best) || isinf_(dr)) return hit;
  uint32_t cur = root;
  for (;;) {
    const srt_bvh_node& n = g.nodes[cur];
    if (n.prim_count == 0) {
      uint32_t L = n.first_child_or_prim_index, R = L + 1;
      bool il, ir;
      float dl = box2(o, d, g.nodes[L], il);
      float dr2 = box2(o, d, g.nodes[R], ir);
      g.cnt->nodes += 2;
      bool pl = dl < best && !isinf_(dl), pr = dr2 < best && !isinf_(dr2);
      if (pl && pr) { bool rf = (il || ir) ? true : !(dl < dr2); st[sp++] = {rf ? L : R, rf ? dl : dr2}; cur = rf ? R : L; continue; }
      if (pl || pr) { cur = pl ? L : R; continue; }
    } else {
      for (uint32_t i = 0; i < n.prim_count; ++i) {
        const srt_triangle& tri = g.triangles[n.first_child_or_prim_index + i];
        V3 tn; g.cnt->tris++;
        if (IntersectsTriangle(o, d, ld3(g.vertices[tri.v0_idx].vertex), ld3(g.vertices[tri.v1_idx].vertex), ld3(g.vertices[tri.v2_idx].vertex), best, tn)) hit = n.first_child_or_prim_index + i;
      }
    }
    for (;;) { if (sp == 0) return hit; --sp; if (st[sp].dist < best) { cur = st[sp].n; break; } }
  }
}
int main(int argc, char** argv) {
  FILE* f = fopen(argv[1], "rb"); char magic[8]; fread(magic, 1, 8, f);
  std::vector<std::vector<char>> b(5);
  for (int i = 0; i < 5; ++i) { uint64_t n; fread(&n, 8, 1, f); b[i].resize(n); fread(b[i].data(), 1, n, f); }
  Counters c; memset(&c, 0, sizeof c);
  Inv g; memset(&g, 0, sizeof g); g.cnt = &c;
  g.bvhs = (srt_bvh*)b[0].data(); g.nodes = (srt_bvh_node*)b[1].data(); g.triangles = (srt_triangle*)b[3].data(); g.vertices = (srt_vertex*)b[4].data();
  std::mt19937 rng(1); std::uniform_real_distribution<float> U(-1, 1);
  int diff_t = 0, diff_id = 0, n = 0, hits = 0; uint64_t rn[2]={0,0}, rt[2]={0,0}, on[2]={0,0}, ot[2]={0,0}, cnt[2]={0,0};
  // primary-like rays, then bounce rays from hit points
  for (int i = 0; i < 400000; ++i) {
    V3 o = v3(0, 9, 40), d = v3(U(rng) * 0.5f, U(rng) * 0.5f, -1.0f);
    for (int bounce = 0; bounce < 3; ++bounce) {
      float b0 = inf_(), b1 = inf_(); V3 tn = v3(0);
      uint64_t n0 = c.nodes, t0 = c.tris;
      uint32_t h0 = Intersects(g, 0, o, d, b0, tn);
      rn[bounce>0] += c.nodes - n0; rt[bounce>0] += c.tris - t0; n0 = c.nodes; t0 = c.tris;
      uint32_t h1 = ordered(g, 0, o, d, b1);
      on[bounce>0] += c.nodes - n0; ot[bounce>0] += c.tris - t0; cnt[bounce>0]++;
      ++n; if (h0 != (uint32_t)-1) ++hits;
      if (f2u(b0) != f2u(b1)) { ++diff_t; if (diff_t < 6) printf("bounce %d: ref tri %u t=%.9g  ordered tri %u t=%.9g\n", bounce, h0, b0, h1, b1); }
      else if (h0 != h1) ++diff_id;
      if (h0 == (uint32_t)-1) break;
      o = (b0 * d) + o; d = normalize_(v3(U(rng), U(rng), U(rng)));
    }
  }
  for (int k=0;k<2;++k) printf("%s: ref nodes/ray %.1f tris/ray %.1f | hybrid nodes/ray %.1f tris/ray %.1f\n", k?"bounce":"primary", (double)rn[k]/cnt[k], (double)rt[k]/cnt[k], (double)on[k]/cnt[k], (double)ot[k]/cnt[k]);
  printf("rays %d hits %d diff_t %d diff_id_only %d\n", n, hits, diff_t, diff_id);
}
