// srt_ctx.h -- the context behind the C-ABI's opaque srt_ctx* (shared by srt_api.cu and srt_multi.cu).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

#include "srt_b200.h"
#include "srt_kernels.h"
#include "srt_finder.h"
#include "srt_relayout.h"

struct srt_group;  // srt_multi.cu: this context's place in a multi-GPU group (NULL = a single GPU)

using namespace srt;

struct srt_ctx {
  int device = 0;
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;
  std::string error;

  // host copies of what the reference keeps in its file-static vectors (gpu_loader.cpp:44-52)
  std::vector<srt_bvh> h_bvhs;
  std::vector<BvhRecord> h_bvh_records;
  uint32_t n_nodes = 0, n_tris = 0, n_mats = 0;
  int tree_depth = 0;
  bool have_scene = false;
  size_t top_pair_bytes = 0;             // breadth-first prefix of d_pairs (top BVH levels)
  cudaStream_t l2_window_stream = nullptr;  // stream the access-policy window is currently set on
  bool l2_window_on = false;
  const void* l2_window_ptr = nullptr;      // what the window covers
  size_t l2_window_bytes = 0;
  size_t l2_limit_set = (size_t)-1;         // last cudaLimitPersistingL2CacheSize this context asked for
  int l2_max_window = -1, l2_max_persist = -1;  // device limits (-1: not asked yet)

  // device scene
  NodePair* d_pairs = nullptr;
  uint2* d_leaves = nullptr;
  TriRecord* d_tris = nullptr;
  MatRecord* d_mats = nullptr;
  BvhRecord* d_bvhs = nullptr;
  // hit finder (srt_finder.h): built at upload unless SRT_OPT_FINDER is 0
  NodePair* d_fpairs = nullptr;
  TriRecord* d_ftris = nullptr;
  RefNode* d_rnodes = nullptr;
  uint32_t* d_tri_leaf = nullptr;
  int finder_mode = 2;           // SRT_OPT_FINDER: 0 off, 1 on, 2 where the sample rays say it pays
  bool finder_ready = false;     // the arrays above describe the uploaded scene
  bool finder_worthwhile = false;
  int finder_depth = 0;
  double finder_cost = 0.0, reference_cost = 0.0;
  // textured-albedo option (SRT_OPT_TEXTURES; include/srt_texmath.h)
  bool textures_on = false;
  TexTri* d_tex_tris = nullptr;            // built by srt_upload_scene when a material says use_texture
  TexInfo* d_tex_info = nullptr;
  unsigned char* d_texels = nullptr;
  std::vector<unsigned char> h_texels;     // RGBA8 of every texture set so far
  std::vector<TexInfo> h_tex_info;
  std::vector<uint32_t> h_mat_tex;         // per material: 1 + texture index, 0 = none
  std::vector<MatRecord> h_mats;
  bool textures_dirty = false;
  float2* d_noise_xy = nullptr;
  float* d_noise_x = nullptr;
  uint32_t noise_count = 0;
  char* noise_stage = nullptr;  // pinned staging buffer of the packed noise tables
  size_t noise_stage_bytes = 0;
  std::map<void**, size_t> dev_capacity;  // bytes allocated behind the scene's device pointers (upload_vec)
  float* d_noise_raw = nullptr; // device copy of the two tables as given (pinned callers: packed on the device)
  size_t noise_raw_bytes = 0;

  // images
  int width = 0, height = 0;
  float4* d_accum = nullptr;
  uchar4* d_out = nullptr;
  uint32_t* d_ids = nullptr;
  bool ids_enabled = false;
  int row0 = 0, row1 = 0;
  int stripe_rank = 0, stripe_world = 1;

  // uniforms + lights
  srt_params params;
  bool have_params = false;
  srt_light lights[SRT_MAX_LIGHTS];
  int light_count = 0;

  // options
  int traversal = TRAV_ORDERED;
  int count_traversal = 0;  // SRT_OPT_COUNT_TRAVERSAL: 0 off, 1 the shader's walk (finder off), 2 what the default path executes
  int l2_persist = 1;  // SRT_OPT_L2_PERSIST: 0 off, 1 automatic, 2 node pairs' top levels, 3 noise tables
  int block_threads = 64;
  int schedule = SCHED_WAVEFRONT;

  // wavefront schedule: path slots + queues (sized for the pixels this context renders), polling state
  PathPool pool;
  uint32_t pool_slots = 0;
  size_t pool_list_cap = 0;
  // A second set of slot records, queues and samples (SRT_OPT_OVERLAP): the chunks of the wavefront schedule alternate
  // between the two, a chunk's init + waves run on wave_stream, its tail kernel and resolve on `stream` -- so the tail of
  // one chunk (a third of a rank's step at 8 GPUs, whatever the rank's share) runs beside the waves of the next one, of the
  // same srt_render call or of the next call.  Images (accum, out) are only ever touched on `stream`, in call order.
  PathPool pool_b;
  uint32_t* d_culled_b = nullptr;
  uint32_t pool_b_slots = 0;
  size_t sample_entries_b = 0;
  int overlap = 1;                               // SRT_OPT_OVERLAP
  int pool_turn = 0;                             // which set the next chunk takes
  cudaStream_t wave_stream = nullptr;
  cudaEvent_t ev_inputs = nullptr;               // recorded on `stream` after uploads; wave_stream waits for it
  cudaEvent_t ev_waves_done[2] = {nullptr, nullptr};
  cudaEvent_t ev_pool_free[2] = {nullptr, nullptr};
  bool pool_busy[2] = {false, false};
  bool inputs_dirty = true;                      // an upload since the last wave_stream launch
  WaveCounters* h_live = nullptr;  // pinned ring the device reports its wave counters into (live slots, paths handed out)
  cudaEvent_t wave_events[4] = {nullptr, nullptr, nullptr, nullptr};
  WaveFork fork = {nullptr, nullptr, nullptr};  // second stream of the finder waves (launch_wf_wave); SRT_NO_FORK=1 in the environment keeps it unused
  bool use_fork = true;
  uint64_t waves = 0;
  uint64_t tail_launches = 0;
  // optional per-kernel timing of the wavefront schedule (SRT_OPT_TIME_KERNELS): three events per wave
  bool time_kernels = false;
  std::vector<cudaEvent_t> timing_events;
  size_t timing_used = 0;
  int refill_threshold = 0;                    // idle lanes that send a trace warp back to the queue (SRT_OPT_REFILL; 0 = each kernel's own default)
  uint32_t wave_slots = (128u << 20) - 32u;     // path slots of the pool (SRT_OPT_WAVE_SLOTS); ray ids (2 per slot) stay below 2^28
  int device_relayout = 2;                      // SRT_OPT_DEVICE_RELAYOUT: 0 host walk, 1 device, 2 device above SRT_FINDER_MAX_TRIS triangles
  bool relayout_on_device = false;              // what the last upload did
  float relayout_device_ms = 0.0f;
  int device_finder_build = 1;                  // SRT_OPT_DEVICE_FINDER_BUILD: the hit finder's trees are built by kernels (srt_finder_build.cu)
  FinderBuildWorkspace finder_ws;
  RelayoutWorkspace relayout_ws;                // raw arrays + scratch of the device re-layout, kept between uploads
  int fused_primary = 0;                        // SRT_OPT_FUSED_PRIMARY: the first wave's find kernel makes the camera rays (no init kernel); measured slower, off
  int sort_rays = 0;                            // SRT_OPT_SORT_RAYS: 0 off (default: measured slower, see srt_b200.h), 1 on
  bool primary_cull = true;                     // SRT_OPT_PRIMARY_CULL
  uint32_t* d_culled = nullptr;                 // bit map of the paths the cull ended (one bit per sample entry)
  uint32_t tail_paths = 1u << 20;               // live paths at which the tail kernel takes over (SRT_OPT_TAIL_PATHS, 0 = never)
  bool tail_paths_set = false;                  // the caller chose it; otherwise scenes with big leaves get a lower one (max_leaf)
  uint32_t max_leaf = 0;                        // most triangles in one leaf of the uploaded reference tree
  size_t sample_entries = 0;                   // float4 entries allocated in pool.samples
  size_t sample_budget_bytes = (size_t)4 << 30;  // sample buffer budget per frame chunk (SRT_OPT_SAMPLE_BUDGET_MB)

  DeviceCounters* d_counters = nullptr;
  uint64_t kernel_launches = 0;

  srt_group* group = nullptr;  // multi-GPU: partition + collective inside srt_render (srt_multi.cu)
};


// srt_api.cu
int srt_fail(srt_ctx* ctx, const char* fmt, ...);
// srt_multi.cu: hooks of the single-context entry points
int srt_group_before_render(srt_ctx* ctx, int32_t n_frames, int32_t* first_frame, int32_t* my_frames);
int srt_group_after_render(srt_ctx* ctx, int32_t total_frames);
void srt_group_release(srt_ctx* ctx);
