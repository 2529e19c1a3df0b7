// srt_device.cuh -- device-side data layout of the B200 path tracer (private to libsrt_b200.so).
//
// The C-ABI takes the reference's std430 arrays (include/srt_types.h); srt_api.cu re-lays them out once
// per upload into the gather-friendly forms below.  Every record is a multiple of 16 B and is read with
// 128-bit __ldg loads (ld.global.nc.v4).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "srt_detmath.h"
#include "srt_types.h"

namespace srt {

// One record per INNER node of the reference BVH (raytrace_types.glsl:41-46), holding the boxes of both
// children so that a single 64 B fetch decides both box tests (the reference fetches the child nodes one at
// a time, ray_intersects.glsl:107).  Child references:
//   bit 31 clear: index of the child's own NodePair record (child is an inner node)
//   bit 31 set  : index into the leaf table (child is a leaf)
// and, for kernels that walk a leaf triangle by triangle until TriRecord's `remaining` mark says 0, the same two
// children as DIRECT references: bit 31 set = index of the leaf's first triangle (no leaf-table fetch on arrival).
struct NodePair {
  float4 a;  // L.min.xyz, L.max.x
  float4 b;  // L.max.yz,  R.min.xy
  float4 c;  // R.min.z,   R.max.xyz
  uint4 d;   // L.ref, R.ref, L.direct, R.direct
};
static_assert(sizeof(NodePair) == 64, "NodePair is one 64 B fetch");

#define SRT_LEAF_BIT 0x80000000u

// One record per triangle in the reference's triangle order (= BVH leaf order, bvh.h:66-72), positions
// pre-gathered through the index record (the reference does 4 dependent gathers per triangle,
// ray_intersects.glsl:114-118).  e1/e2 are the IEEE differences v1-v0, v2-v0 the shader recomputes per test
// (ray_intersects.glsl:69-70) -- same bits, computed once.
struct TriRecord {
  float4 v0;  // v0.xyz, bits: material index
  float4 e1;  // edge_1.xyz, bits: rank in the reference's visiting order (tie-break key)
  float4 e2;  // edge_2.xyz, bits: triangles that follow this one in its leaf (0 marks the leaf's last triangle)
};
static_assert(sizeof(TriRecord) == 48, "TriRecord is three 16 B fetches");

// MaterialFromOBJ (raytrace_types.glsl:19-27) after TriangleToSupportedMat (raytrace_utils.glsl:140-175).
struct MatRecord {
  float4 a;  // albedo (= Kd, decree Q7), roughness = 1/(Ns+1e-7)
  float4 b;  // specular (= Ks), bits: 1 + index of the material's texture (0 = none; textured-albedo option)
};

// BVH (raytrace_types.glsl:32-38) plus the root's own box (the reference box-tests the root when it pops
// it, ray_intersects.glsl:102-110).
struct BvhRecord {
  float frame[16];
  float root_min[3];
  uint32_t root_ref;
  float root_max[3];
  uint32_t root_direct;  // root as a direct reference (see NodePair::d)
  // hit finder (srt_finder.h): the root of this BVH's private tree as a direct reference into DeviceScene::fpairs /
  // ftris, the reference node the BVH starts at (where a replay stops), the finder's box padding
  uint32_t finder_root;
  uint32_t root_node;
  float finder_pad;
  uint32_t _pad;
};
static_assert(sizeof(BvhRecord) == 112, "BvhRecord: frame first (srt_update_frame rewrites its 64 bytes)");

// Hit finder (srt_finder.h, srt_wf_find.cuh).  One record per node of the reference BVH in the reference's own
// numbering: the box the shader tests when it pops the node (ray_intersects.glsl:107-110), the node above it and the
// visiting rank of the first triangle below it -- what a replay of the shader's culls needs.
struct RefNode {
  float4 a;  // min.xyz, bits: parent node (0xFFFFFFFF above a root)
  float4 b;  // max.xyz, bits: rank of the first triangle the shader's traversal meets in this subtree
};
#define SRT_CAND_MAX 8          // candidates kept per ray; a ray with more is traced by wf_trace_kernel instead
#define SRT_CAND_BVH_SHIFT 26   // candidate word: triangle | bvh << 26
#define SRT_FINDER_MAX_TRIS (1u << 21)
#define SRT_FINDER_DEVICE_MIN_TRIS (1u << 14)  // below this the host builds the finder's trees faster than ~20 levels of launches (Rubik: 0.3 vs 1.6 ms)

// SRT_OPT_TRAVERSAL values (srt_b200.h)
enum { TRAV_REFERENCE = 0, TRAV_ORDERED = 1 };
// SRT_OPT_SCHEDULE values (srt_b200.h)
enum { SCHED_PER_PATH = 0, SCHED_REGEN = 1, SCHED_PERSISTENT = 2, SCHED_WAVEFRONT = 3 };

#define SRT_MAX_BVHS 64

// Textured-albedo option (SRT_OPT_TEXTURES, include/srt_texmath.h): per triangle the three texture coordinates and
// the BVH the triangle belongs to (its frame gives the model-space hit point the shader interpolates at,
// raytrace_compute.glsl:156-159); per texture where its RGBA8 texels start.
struct TexTri {
  float uv[6];
  uint32_t bvh;
  uint32_t _pad;
};
struct TexInfo {
  uint32_t offset;  // first texel in DeviceScene::texels
  uint32_t width, height, _pad;
};

struct DeviceScene {
  const TexTri* tex_tris;        // null unless the option is on AND some material has a texture
  const TexInfo* tex_info;
  const unsigned char* texels;   // RGBA8
  const NodePair* pairs;
  const uint2* leaves;     // (first triangle, count) per leaf
  const TriRecord* tris;
  const MatRecord* mats;
  const BvhRecord* bvhs;
  // hit finder, null when the scene has none: a private BVH2 over the same triangles (padded boxes, leaves of <= 4,
  // its own copy of the triangle records in ITS leaf order: v0.w = the triangle's index in `tris`), the reference
  // nodes for the replay, and the reference leaf (node index) of every triangle
  const NodePair* fpairs;
  const TriRecord* ftris;
  const RefNode* rnodes;
  const uint32_t* tri_leaf;
  const float2* noise_unit_xy;  // noiseTex.xy   (raytrace_utils.glsl:15)
  const float* noise_uniform_x; // noiseUniformTex.x (raytrace_utils.glsl:52)
};

struct DeviceLight {
  float px, py, pz, intensity;
  float cx, cy, cz, pad;
};

// Everything a frame needs that is not a big array: the uniform block (raytrace_compute.glsl:18-31,39), the
// camera GetCamera derives from it (raytrace_compute.glsl:47-76, hoisted to the host: same IEEE ops, once per
// launch instead of once per pixel) and the padded light list (decree Q2).
struct FrameParams {
  int width, height;
  int n_texels;          // Width*Height
  float width_f, height_f;
  int accum_frames;      // accumFrames of the first frame of this launch
  int n_frames;          // frames rendered by this launch
  int y0, y1;            // row range
  int stripe_rank, stripe_world;  // of the 8-row bands inside [y0,y1) this launch renders those with band % world == rank
  int light_count;
  float light_count_f;
  uint32_t bvh_count;
  int max_depth;
  int stack_cap;         // entries per thread in the shared-memory traversal stack
  V3 center, pixel00, delta_u, delta_v;
  DeviceLight lights[SRT_MAX_LIGHTS];
};

struct DeviceCounters {
  unsigned long long segments, shadow_rays, paths, capped_paths, nodes, tris;
  unsigned long long culled;  // camera rays ended by the primary cull (counted with SRT_OPT_COUNT_TRAVERSAL only)
  unsigned int next_tile;  // work counter of the persistent kernel (zeroed before each launch)
  unsigned int check_flags;  // checked build (-DSRT_CHECKED) only: one bit per kind of bounds violation seen (SRT_CHK_*)
};

// Checked build (make libsrt_b200_checked.so): compute-sanitizer is not available on this pool, so the kernels of the
// wavefront schedule carry their own bounds checks -- every push on a shared-memory traversal stack, every index into a
// queue / list / slot array / sample buffer.  A violation sets its bit in DeviceCounters::check_flags (read through
// srt_get_counters) and the access is still made, as in the normal build.  The normal build compiles the checks away.
#define SRT_CHK_STACK 0      // traversal stack deeper than fp.stack_cap
#define SRT_CHK_LIST_READ 1  // queue entry read beyond pool.list_cap
#define SRT_CHK_SLOT 2       // slot index beyond pool.cap
#define SRT_CHK_LIST_WRITE 3 // queue entry written beyond pool.list_cap
#define SRT_CHK_STASH 4      // logic kernel's put-aside list beyond its 64 entries
#define SRT_CHK_SAMPLE 5     // sample written beyond the launch's paths
#define SRT_CHK_CAND 6       // candidate index beyond SRT_CAND_MAX / replay position beyond the slots
#define SRT_CHK_OVERFLOW 7   // overflow list written beyond pool.list_cap
#define SRT_CHK_PIXEL 8      // pixel / texel index beyond the image
#ifdef SRT_CHECKED
#define SRT_CHECK(cond, bit)                                        \
  do {                                                              \
    if (!(cond)) atomicOr(&counters->check_flags, 1u << (bit));     \
  } while (0)
#else
#define SRT_CHECK(cond, bit) ((void)0)
#endif

// ---- wavefront schedule (srt_wavefront.cuh): queues, path slots, per-frame sample buffer ---------------------------
struct WaveCounters {
  unsigned int n_closest[2];  // entries at the front of list[parity]
  unsigned int n_shadow[2];   // entries at the back of list[parity]
  unsigned int cursor;        // fetch cursor of wf_trace_kernel
  unsigned int live_slots;    // slots that still carry a path
  unsigned long long next_path;  // next path (frame-major: frame * n_pixels + pixel slot) to hand to a free slot
  unsigned int n_dying[2];    // entries of dlist[parity]
  unsigned int cursor2;       // fetch cursor of a wf_trace_kernel launch that takes the shadow queue alone
  unsigned int _pad;
};

// Queues between wf_find_kernel and its two consumers (srt_wf_find.cuh); zeroed by wf_init_kernel / wf_logic_kernel.
struct FinderQueues {
  WaveCounters over;      // rays with more than SRT_CAND_MAX candidates: n_closest[0] entries of PathPool::olist, traced
                          // by wf_trace_kernel (given a PathPool whose q / list[0] point here)
  unsigned int n_replay;  // entries of PathPool::rlist
  unsigned int cursor;    // fetch cursor of wf_find_kernel
  unsigned int _pad[2];
};

// Per-slot records (SoA).  A slot carries ONE path at a time; when the path ends the slot takes the next path of the
// launch from WaveCounters::next_path.  A path has up to TWO rays in flight per wave: the closest-hit ray of its next
// segment and the shadow ray of the surface it has just left (the direct-light term that ray decides is computed when
// the ray is emitted and added when its answer is in: `pend`).  What the queues and the trace kernels handle are RAY IDS,
//   ray id = 2 * slot (the slot's closest-hit ray) or 2 * slot + 1 (its shadow ray),
// which index `ray` and `hit`; to those kernels a ray id is an opaque record number.
struct PathPool {
  // records that are always read together share one 32 B sector (they are gathered through the queues)
  float4* ray;   // [2*id]   ray origin.xyz, tmax (initial intersection_distance)
                 // [2*id+1] ray direction.xyz, bits: kind (SRT_WF_KIND_*)
  uint2* hit;    // [id] result: primitive (or SRT_NO_HIT), distance bits.  hit[2*slot+1].x = SRT_WF_NOT_PENDING while the
                 // slot has no shadow ray in flight (written by whoever decides not to emit one; not looked at while the
                 // path has not bounced yet, iter == 0: it cannot have one)
  float4* pst;   // [2*slot]   throughput.xyz, bits: the path's number in the launch (frame * n_pixels + pixel slot, < 2^32)
                 // [2*slot+1] colour.xyz, bits: depth | rand_index << 8 | iter << 20
  float4* pend;  // while a shadow ray is out: the term `throughput * SampleDirect * weight` for an unoccluded light (.xyz), and
                 // in .w two bits per channel saying what the occluded term is (+0, -0 or NaN: 0 * anything)
  uint32_t* dlist[2];  // cap entries each: SLOTS whose path ends as soon as its last shadow ray is answered (no closest-hit
                       // ray of theirs is in the wave, so the closest queue does not lead the logic kernel to them)
  uint32_t* list[2];   // list_cap entries each: closest-hit queue from the front, shadow queue from the back
  uint32_t cap;        // slots
  uint32_t ray_cap;    // ray ids: 2 * cap
  uint32_t list_cap;   // 2 * cap + room for the holes the logic warps' chunked reservations can leave
  WaveCounters* q;
  // hit finder only (null otherwise): per ENTRY OF THE REPLAY LIST SRT_CAND_MAX candidates (triangle | bvh << 26, distance bits), the
  // list of rays to replay (slot | candidates << 28, theta bits), the list of rays to trace the long way, their counters
  uint2* cand;
  uint2* rlist;
  uint32_t* olist;
  FinderQueues* fq;
  // one fp32 sample per path of the launch, frame-major, summed per pixel in frame order by wf_resolve_kernel
  float4* samples;
  // primary cull only (null otherwise): one bit per path of the launch, frame-major like `samples`; a set bit says the
  // path ended in wf_init_kernel with the sky sample, which is then NOT stored (wf_resolve_kernel adds the constant)
  uint32_t* culled;
  // SRT_OPT_SORT_RAYS only (null otherwise; srt_wf_sort.cuh): the bins of the counting sort, per queue entry its key, and
  // the sorted copy of the queue
  uint32_t* sort_bins;
  uint32_t* sort_keys;
  uint32_t* sort_tmp;
  uint32_t n_pixels;           // pixel slots of this launch (8x4 tiles x 32)
  unsigned long long n_paths;  // n_pixels * frames of this launch
};

}  // namespace srt
