// srt_multi.cu -- several B200s behind one srt_render: partition of the frame + ONE collective per call, inside the library.
//
// No reference call site: src/main.cpp:706 is a single glDispatchCompute on one GPU.  SURVEY.md 8(e): pixels are
// independent (each invocation RMWs only its own accumulation texel, raytrace_compute.glsl:388,404-406) and samples are
// independent given accumFrames (it enters through `samp`, :400-401, and the final divide, :412), so the frame shards
// both ways with full replicas of scene + noise on every GPU and no exchange inside the path loop:
//   SRT_SPLIT_BANDS   (default) rank r renders the interleaved 8-row bands b = r (mod world) of every frame.  After the
//                     render each rank packs ITS rows (P*16/world bytes) and sends them to rank 0, which scatters them
//                     into the gathered accumulation image: the bits of a single-GPU render.
//   SRT_SPLIT_SAMPLES rank r renders the whole image for its share of the call's frames (accumFrames .. +n split into
//                     `world` consecutive ranges); one ncclReduce(sum) of the fp32 accumulation images into rank 0's
//                     gathered image.  fp32 additions are re-ordered: equal to the single-GPU image within tolerance
//                     (tests: 1e-4 relative), not in bits.
// Every rank's own accumulation image only ever holds what that rank rendered, and rank 0 tonemaps from the GATHERED
// image, so progressive rendering (render, render, ... without a reset) never counts a contribution twice.
//
// Two ways in:
//   srt_group_join     one process per GPU (torchrun / mpirun): the launcher hands every rank the 128-byte id that rank 0
//                      got from srt_group_unique_id (bench.py broadcasts it with torch.distributed; tests with a file)
//   srt_create_multi   one process, n GPUs: returns a HEAD context; uploads / uniforms / resize fan out to a private
//                      context per device, srt_render runs them on one host thread each, reads come from rank 0
// NCCL is loaded with dlopen when a group is first formed: libsrt_b200.so has no link-time dependency on it, and a
// process that already holds a libnccl.so.2 (torch's) shares that copy.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>  // types only; every call goes through the table below
#include <stdio.h>
#include <string.h>

#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#include "srt_ctx.h"

namespace {

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::string error;
};

NcclApi& nccl() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
      api.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
      if (api.lib) break;
    }
    if (!api.lib) {
      api.error = std::string("NCCL not found (dlopen libnccl.so.2): ") + (dlerror() ? dlerror() : "");
      return;
    }
    auto sym = [&](const char* n) {
      void* p = dlsym(api.lib, n);
      if (!p && api.error.empty()) api.error = std::string("NCCL symbol missing: ") + n;
      return p;
    };
    api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
    api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
    api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
    api.Reduce = (decltype(api.Reduce))sym("ncclReduce");
    api.Send = (decltype(api.Send))sym("ncclSend");
    api.Recv = (decltype(api.Recv))sym("ncclRecv");
    api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
    api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
  });
  return api;
}

// rows of the interleaved 8-row bands a rank owns -> a contiguous buffer and back (float4 per texel)
__global__ void pack_bands_kernel(const float4* __restrict__ image, float4* __restrict__ packed, int width, int height, int rank,
                                  int world, int unpack_into_image) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;  // texel of the packed buffer
  const int n_bands = height / 8;
  const int my_bands = (n_bands - rank + world - 1) / world;
  const size_t n = (size_t)my_bands * 8u * (size_t)width;
  if (i >= n) return;
  const size_t row = i / (size_t)width, x = i - row * (size_t)width;
  const size_t band = (size_t)rank + (row / 8u) * (size_t)world;
  const size_t y = band * 8u + (row & 7u);
  if (unpack_into_image) const_cast<float4*>(image)[y * (size_t)width + x] = packed[i];
  else packed[i] = image[y * (size_t)width + x];
}

// the summed images carry alpha = world; the accumulation image's alpha is 1 (raytrace_compute.glsl:406)
__global__ void alpha_one_kernel(float4* __restrict__ image, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) image[i].w = 1.0f;
}

size_t band_texels(int width, int height, int rank, int world) {
  const int n_bands = height / 8;
  const int my_bands = (n_bands - rank + world - 1) / world;
  return (size_t)my_bands * 8u * (size_t)width;
}

}  // namespace

struct srt_group {
  int rank = 0, world = 1;
  ncclComm_t comm = nullptr;
  int split = SRT_SPLIT_BANDS;
  // rank 0: the gathered accumulation image (what srt_read_accum / srt_tonemap / srt_accum_device_ptr see)
  float4* d_gathered = nullptr;
  size_t gathered_texels = 0;
  float4* d_pack = nullptr;   // this rank's packed rows (send side), or rank 0's staging for every rank (receive side)
  size_t pack_texels = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending;  // collectives whose duration has not been read yet
  double collective_ms = 0.0;
  uint64_t collectives = 0;
  int32_t call_first = 0, call_frames = 0;  // the frames of the call in flight (set by before_render)
  // srt_create_multi: the head owns one context per device (peers[0] is the head itself)
  std::vector<srt_ctx*> peers;
  bool fanning = false;
};

#define GCU(ctx, call)                                                                          \
  do {                                                                                          \
    cudaError_t e_ = (call);                                                                    \
    if (e_ != cudaSuccess) return srt_fail(ctx, "%s: %s", #call, cudaGetErrorString(e_));       \
  } while (0)
#define GNC(ctx, call)                                                                          \
  do {                                                                                          \
    ncclResult_t r_ = (call);                                                                   \
    if (r_ != ncclSuccess) return srt_fail(ctx, "%s: %s", #call, nccl().GetErrorString(r_));    \
  } while (0)

static int group_attach(srt_ctx* ctx, int rank, int world, ncclComm_t comm) {
  srt_group* g = new srt_group();
  g->rank = rank;
  g->world = world;
  g->comm = comm;
  ctx->group = g;
  ctx->stripe_rank = rank;
  ctx->stripe_world = world;
  return 0;
}

extern "C" int srt_group_unique_id(void* id, uint64_t bytes) {
  if (!id || bytes < sizeof(ncclUniqueId)) return srt_fail(nullptr, "srt_group_unique_id: need %zu bytes", sizeof(ncclUniqueId));
  NcclApi& api = nccl();
  if (!api.error.empty()) return srt_fail(nullptr, "%s", api.error.c_str());
  ncclUniqueId u;
  ncclResult_t r = api.GetUniqueId(&u);
  if (r != ncclSuccess) return srt_fail(nullptr, "ncclGetUniqueId: %s", api.GetErrorString(r));
  memset(id, 0, bytes);
  memcpy(id, &u, sizeof u);
  return 0;
}

extern "C" int srt_group_join(srt_ctx* ctx, int32_t rank, int32_t world, const void* id, uint64_t bytes) {
  if (!ctx) return 1;
  if (ctx->group) return srt_fail(ctx, "srt_group_join: this context is in a group already");
  if (world < 1 || rank < 0 || rank >= world) return srt_fail(ctx, "srt_group_join: rank %d of world %d", rank, world);
  if (!id || bytes < sizeof(ncclUniqueId)) return srt_fail(ctx, "srt_group_join: need the %zu-byte id of srt_group_unique_id", sizeof(ncclUniqueId));
  NcclApi& api = nccl();
  if (!api.error.empty()) return srt_fail(ctx, "%s", api.error.c_str());
  GCU(ctx, cudaSetDevice(ctx->device));
  ncclUniqueId u;
  memcpy(&u, id, sizeof u);
  ncclComm_t comm = nullptr;
  GNC(ctx, api.CommInitRank(&comm, world, u, rank));
  return group_attach(ctx, rank, world, comm);
}

void srt_group_release(srt_ctx* ctx) {
  srt_group* g = ctx->group;
  if (!g) return;
  cudaSetDevice(ctx->device);
  if (g->comm) nccl().CommDestroy(g->comm);
  if (g->d_gathered) cudaFree(g->d_gathered);
  if (g->d_pack) cudaFree(g->d_pack);
  for (auto& p : g->pending) { cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
  for (size_t i = 1; i < g->peers.size(); ++i) srt_destroy(g->peers[i]);
  delete g;
  ctx->group = nullptr;
}

// the image rank 0 shows: the gathered one in a group, the context's own otherwise
float4* srt_group_display_accum(srt_ctx* ctx) {
  return (ctx->group && ctx->group->rank == 0 && ctx->group->world > 1 && ctx->group->d_gathered) ? ctx->group->d_gathered : ctx->d_accum;
}

// Before the kernels of a call are queued: which of the call's frames this rank renders.
int srt_group_before_render(srt_ctx* ctx, int32_t n_frames, int32_t* first_frame, int32_t* my_frames) {
  srt_group* g = ctx->group;
  *first_frame = ctx->params.accum_frames;
  *my_frames = n_frames;
  if (!g || g->world == 1) return 0;
  g->call_first = ctx->params.accum_frames;
  g->call_frames = n_frames;
  if (g->split == SRT_SPLIT_SAMPLES) {
    // consecutive ranges with edges first + floor(n * r / world) (frame_split.py:sample_ranges is the same rule)
    const long long n = n_frames, w = g->world, r = g->rank;
    const int before = (int)((n * r) / w);
    *first_frame = ctx->params.accum_frames + before;
    *my_frames = (int)((n * (r + 1)) / w) - before;
    ctx->stripe_rank = 0;
    ctx->stripe_world = 1;
  } else {
    ctx->stripe_rank = g->rank;
    ctx->stripe_world = g->world;
  }
  return 0;
}

// After them, on the same stream: the call's ONE collective, then rank 0's tonemap of the gathered image.
// total_frames = accumFrames of the last frame of the call (the divisor of raytrace_compute.glsl:412).
int srt_group_after_render(srt_ctx* ctx, int32_t total_frames) {
  srt_group* g = ctx->group;
  if (!g || g->world == 1) return 0;
  NcclApi& api = nccl();
  const size_t texels = (size_t)ctx->width * ctx->height;
  GCU(ctx, cudaSetDevice(ctx->device));
  if (g->rank == 0 && g->gathered_texels != texels) {
    GCU(ctx, cudaStreamSynchronize(ctx->stream));
    if (g->d_gathered) cudaFree(g->d_gathered);
    g->d_gathered = nullptr;
    GCU(ctx, cudaMalloc(&g->d_gathered, texels * sizeof(float4)));
    g->gathered_texels = texels;
  }
  cudaEvent_t e0, e1;
  GCU(ctx, cudaEventCreate(&e0));
  GCU(ctx, cudaEventCreate(&e1));
  GCU(ctx, cudaEventRecord(e0, ctx->stream));
  if (g->split == SRT_SPLIT_SAMPLES) {
    GNC(ctx, api.Reduce(ctx->d_accum, g->rank == 0 ? (void*)g->d_gathered : (void*)ctx->d_accum, texels * 4, ncclFloat, ncclSum, 0,
                        g->comm, ctx->stream));
    ctx->kernel_launches++;
    if (g->rank == 0) {
      alpha_one_kernel<<<(unsigned)((texels + 255) / 256), 256, 0, ctx->stream>>>(g->d_gathered, texels);
      ctx->kernel_launches++;
    }
  } else {
    // every rank's rows, packed: rank 0 receives them side by side and scatters each block into the gathered image
    size_t need = 0, offset[64];
    if (g->world > 64) return srt_fail(ctx, "srt_render: at most 64 ranks");
    if (g->rank == 0) {
      for (int r = 0; r < g->world; ++r) { offset[r] = need; need += band_texels(ctx->width, ctx->height, r, g->world); }
    } else {
      need = band_texels(ctx->width, ctx->height, g->rank, g->world);
    }
    if (g->pack_texels < need) {
      GCU(ctx, cudaStreamSynchronize(ctx->stream));
      if (g->d_pack) cudaFree(g->d_pack);
      g->d_pack = nullptr;
      GCU(ctx, cudaMalloc(&g->d_pack, (need ? need : 1) * sizeof(float4)));
      g->pack_texels = need;
    }
    const size_t mine = band_texels(ctx->width, ctx->height, g->rank, g->world);
    const unsigned blocks = (unsigned)((mine + 255) / 256);
    if (g->rank != 0) {
      if (mine) pack_bands_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->d_accum, g->d_pack, ctx->width, ctx->height, g->rank, g->world, 0);
      GNC(ctx, api.Send(g->d_pack, mine * 4, ncclFloat, 0, g->comm, ctx->stream));
      ctx->kernel_launches += 2;
    } else {
      if (mine) pack_bands_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->d_accum, g->d_pack, ctx->width, ctx->height, 0, g->world, 0);
      GNC(ctx, api.GroupStart());
      for (int r = 1; r < g->world; ++r) {
        const size_t n = band_texels(ctx->width, ctx->height, r, g->world);
        GNC(ctx, api.Recv(g->d_pack + offset[r], n * 4, ncclFloat, r, g->comm, ctx->stream));
      }
      GNC(ctx, api.GroupEnd());
      for (int r = 0; r < g->world; ++r) {
        const size_t n = band_texels(ctx->width, ctx->height, r, g->world);
        if (n) pack_bands_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(g->d_gathered, g->d_pack + offset[r], ctx->width, ctx->height, r, g->world, 1);
      }
      ctx->kernel_launches += 2 + (uint64_t)g->world;
    }
    GCU(ctx, cudaGetLastError());
  }
  GCU(ctx, cudaEventRecord(e1, ctx->stream));
  g->pending.push_back({e0, e1});
  g->collectives++;
  if (g->rank == 0) {
    GCU(ctx, launch_tonemap(g->d_gathered, ctx->d_out, texels, (float)total_frames, ctx->stream));
    ctx->kernel_launches++;
  }
  return 0;
}

extern "C" int srt_get_collective_ms(srt_ctx* ctx, double* ms, uint64_t* calls) {
  if (!ctx) return 1;
  srt_group* g = ctx->group;
  if (ms) *ms = 0.0;
  if (calls) *calls = 0;
  if (!g) return 0;
  GCU(ctx, cudaSetDevice(ctx->device));
  GCU(ctx, cudaStreamSynchronize(ctx->stream));
  for (auto& p : g->pending) {
    float t = 0.0f;
    cudaEventElapsedTime(&t, p.first, p.second);
    g->collective_ms += t;
    cudaEventDestroy(p.first);
    cudaEventDestroy(p.second);
  }
  g->pending.clear();
  if (ms) *ms = g->collective_ms;
  if (calls) *calls = g->collectives;
  g->collective_ms = 0.0;
  g->collectives = 0;
  return 0;
}

int srt_group_set_split(srt_ctx* ctx, int value) {
  if (value != SRT_SPLIT_BANDS && value != SRT_SPLIT_SAMPLES) return srt_fail(ctx, "SRT_OPT_SPLIT: 0 (image bands) or 1 (sample ranges)");
  if (!ctx->group) return srt_fail(ctx, "SRT_OPT_SPLIT: the context is not in a group (srt_group_join / srt_create_multi)");
  ctx->group->split = value;
  return 0;
}

// ---- one process, n GPUs -------------------------------------------------------------------------------------------------
extern "C" int srt_create_multi(srt_ctx** out, const int* devices, int32_t n) {
  if (!out) return srt_fail(nullptr, "srt_create_multi: out is NULL");
  *out = nullptr;
  if (!devices || n < 1 || n > 64) return srt_fail(nullptr, "srt_create_multi: 1..64 devices");
  NcclApi& api = nccl();
  if (n > 1 && !api.error.empty()) return srt_fail(nullptr, "%s", api.error.c_str());
  std::vector<srt_ctx*> peers((size_t)n, nullptr);
  for (int i = 0; i < n; ++i) {
    if (srt_create(&peers[i], devices[i]) != 0) {
      for (int k = 0; k < i; ++k) srt_destroy(peers[k]);
      return 1;  // srt_last_error(NULL) holds srt_create's message
    }
  }
  if (n > 1) {
    std::vector<ncclComm_t> comms((size_t)n, nullptr);
    ncclResult_t r = api.CommInitAll(comms.data(), n, devices);
    if (r != ncclSuccess) {
      for (srt_ctx* c : peers) srt_destroy(c);
      return srt_fail(nullptr, "ncclCommInitAll: %s", api.GetErrorString(r));
    }
    for (int i = 0; i < n; ++i) group_attach(peers[i], i, n, comms[i]);
    peers[0]->group->peers = peers;
  }
  *out = peers[0];
  return 0;
}

// Fan-out of an entry point over the devices of a head context.  Returns -1 when `ctx` is no head (the caller goes on with
// its own work), else the call's status.  parallel: one host thread per device (srt_render's wave loop waits on events).
int srt_group_fanout(srt_ctx* ctx, bool parallel, const std::function<int(srt_ctx*)>& call) {
  srt_group* g = ctx ? ctx->group : nullptr;
  if (!g || g->peers.size() < 2 || g->fanning) return -1;
  g->fanning = true;
  std::vector<int> rc(g->peers.size(), 0);
  if (parallel) {
    std::vector<std::thread> pool;
    for (size_t i = 1; i < g->peers.size(); ++i) pool.emplace_back([&, i] { rc[i] = call(g->peers[i]); });
    rc[0] = call(ctx);
    for (auto& t : pool) t.join();
  } else {
    for (size_t i = 0; i < g->peers.size(); ++i) rc[i] = call(g->peers[i]);
  }
  g->fanning = false;
  for (size_t i = 0; i < rc.size(); ++i) {
    if (rc[i] != 0) {
      if (i != 0) ctx->error = "device " + std::to_string(g->peers[i]->device) + ": " + g->peers[i]->error;
      return 1;
    }
  }
  return 0;
}

bool srt_group_is_sample_split(const srt_ctx* ctx) { return ctx->group && ctx->group->world > 1 && ctx->group->split == SRT_SPLIT_SAMPLES; }
int srt_group_rank(const srt_ctx* ctx) { return ctx->group ? ctx->group->rank : 0; }
int srt_group_peer_count(const srt_ctx* ctx) { return (ctx && ctx->group && !ctx->group->peers.empty()) ? (int)ctx->group->peers.size() : 1; }
srt_ctx* srt_group_peer(const srt_ctx* ctx, int i) { return ctx->group->peers[(size_t)i]; }
