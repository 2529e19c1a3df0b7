/* srt_b200.h -- C-ABI of the B200 path-tracing hot path (libsrt_b200.so).
 *
 * The reference has no FFI layer: its boundary around the hot path is the set of OpenGL calls the host
 * makes around glDispatchCompute.  Each entry point below replaces one of those call
 * sites (cited as file:line in /root/reference) and takes exactly the host data that call site hands
 * to GL -- plain pointers and sizes in the layouts of srt_types.h, no CUDA or torch types.
 *
 * Conventions: every function returns 0 on success, non-zero on error (srt_last_error gives the
 * text).  Host arrays stay owned by the caller and are copied; device memory is owned by the context
 * and released by srt_destroy.  Calls on one context must be serialised by the caller (the reference is
 * single-threaded on its GL context thread).  There is NO CPU fallback: without a usable CUDA device
 * srt_create fails.
 */
#ifndef SRT_B200_H_
#define SRT_B200_H_

#include "srt_types.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct srt_ctx srt_ctx;

/* Counters of the work done by srt_render since the last srt_reset_counters (device-side counts). */
typedef struct srt_counters {
  uint64_t segments;     /* CheckHit calls of the path loop (raytrace_compute.glsl:225): primary + bounce rays */
  uint64_t shadow_rays;  /* CheckHit calls of CheckLightOccluded (raytrace_compute.glsl:174) */
  uint64_t paths;        /* pixel samples started */
  uint64_t capped_paths; /* paths cut by the 1024-iteration decree (Q9) */
  uint64_t kernel_launches; /* kernels launched by this context (all kinds) */
  uint64_t nodes;        /* BVH nodes box-tested   (only counted when SRT_FLAG_COUNT_TRAVERSAL is set) */
  uint64_t tris;         /* triangles tested       (only counted when SRT_FLAG_COUNT_TRAVERSAL is set) */
  uint64_t culled_primaries; /* camera rays that missed every root box and ended where they were generated
                                (SRT_OPT_PRIMARY_CULL; only counted when SRT_OPT_COUNT_TRAVERSAL is set) */
  uint64_t check_flags;  /* always 0 in the normal build; in the checked build (libsrt_b200_checked.so, -DSRT_CHECKED) one
                            bit per kind of out-of-bounds access the kernels' own checks saw (csrc/srt_device.cuh SRT_CHK_*) */
} srt_counters;

/* srt_set_option keys */
#define SRT_OPT_TRAVERSAL 1      /* 0 = the shader's visiting order everywhere (bit-exact by construction)
                                    1 = nearer child first where neither child's box contains the ray origin, the
                                        shader's order elsewhere (default; same hits bar documented grazing ties) */
#define SRT_OPT_COUNT_TRAVERSAL 2 /* for the roofline figures (slower; the tail kernel is not used while counting):
                                     1 = count the box / triangle tests of the SHADER's walk (every ray through the
                                         reference BVH in its visiting order; the hit finder is switched off)
                                     2 = count the box / triangle tests the DEFAULT path really executes (hit finder +
                                         replay + overflow walk, primary cull) */
#define SRT_OPT_L2_PERSIST 3     /* which array an L2 access-policy window (persisting hits) is put on for the context's stream:
                                    0 = none, 1 (default) and 2 = the BVH's top levels, 3 = the two noise tables (12 B per pixel, read at
                                    hashed positions ~10 times per shaded path).  Measured: no gain from either; 3 is slower on Rubik */
#define SRT_OPT_BLOCK_THREADS 4  /* threads per block of the path kernel (multiple of 32) */
#define SRT_OPT_SCHEDULE 5       /* 0 = one path at a time per lane (the shader's schedule)
                                    1 = per-lane state machine with path regeneration + warp-synchronous traversal
                                    2 = persistent warps over 8x4 tiles, traversal state kept across refills of idle
                                        lanes
                                    3 = wavefront: one path slot per pixel, a persistent traversal kernel fed from ray
                                        queues + full-warp shading kernels, queued wave by wave (default for models)
                                    all four give the same bits; only the lane schedule differs */

#define SRT_OPT_WAVE_SLOTS 6      /* wavefront schedule: path slots in flight, 104 B each (default 128 Mi, never more than the launch has paths) */
#define SRT_OPT_SAMPLE_BUDGET_MB 7 /* wavefront schedule: size limit of the per-path sample buffer; longer launches are cut
                                      into frame chunks (default 4096 MB) */

#define SRT_OPT_REFILL 8          /* wavefront schedule: idle lanes per warp that trigger a queue fetch (1..32; default: 12 in the
                                     trace kernel, 16 in the find kernel) */

#define SRT_OPT_TIME_KERNELS 9    /* wavefront schedule: 1 = bracket every trace / logic launch with CUDA events on the launching
                                     stream (read with srt_get_kernel_times) */

#define SRT_OPT_TEXTURES 10       /* 1 = materials with use_texture sample the image given by srt_set_texture at the interpolated
                                     texture coordinate (raytrace_utils.glsl:145-167 with real handles and coordinates; arithmetic in
                                     srt_texmath.h).  0 (default) = the decree for the reference as shipped: albedo = Kd (Q7).
                                     Schedules 0 and 3 only. */

#define SRT_OPT_TAIL_PATHS 11     /* wavefront schedule: once no more than this many paths are alive (and none is left to start),
                                     ONE kernel runs them all to their last bounce instead of ~50 ever smaller waves; the same
                                     per-path arithmetic, the same bits (default 1 Mi -- 128 Ki for a scene whose reference tree has
                                     leaves of more than 64 triangles, which the tail's lock-step walk is slow on, unless the
                                     option is set; 0 = waves to the end).  Not used with SRT_OPT_TEXTURES or SRT_OPT_COUNT_TRAVERSAL. */

#define SRT_OPT_FINDER 12         /* wavefront schedule: how Intersects (ray_intersects.glsl:99-133) is evaluated.
                                     0 = walk the reference BVH as the shader does
                                     1 = hit finder: collect the triangles the ray passes the shader's triangle test against
                                         through a private SAH tree, then replay the shader's box culls above those few
                                         triangles in its visiting order -- the same hit and distance bit for bit (that of
                                         SRT_OPT_TRAVERSAL = 0), a fraction of the tests
                                     2 = (default) 1 when sample rays at srt_upload_scene say it is less work (Rubik, the
                                         airplane), 0 otherwise (the random soup: every ray meets dozens of triangles)
                                     Set before srt_upload_scene.  Scenes above 2 Mi triangles never use it. */

#define SRT_OPT_PRIMARY_CULL 13   /* wavefront schedule: 1 (default) = a primary ray that misses the root box of every BVH -- the
                                     first test of the shader's own traversal (ray_intersects.glsl:106-110) -- ends where it is
                                     generated: its path's sample is the sky term (raytrace_compute.glsl:287-292), it takes no
                                     path slot and no queue entry.  Applies when every path of a frame chunk starts in wave 0
                                     (chunks are sized for that when the pool holds at least one frame).  0 = every primary ray
                                     goes through the queues.  Same bits either way. */

/* Graphics::Compute::Compute(path) + Init()  (src/graphics/Shader.cpp:14-26,125-157): create the "program".
 * Kernels are AOT-compiled for sm_100a, so there is no source path.  device = CUDA ordinal. */
int srt_create(srt_ctx** out, int device);
/* glDeleteProgram / end of process. */
void srt_destroy(srt_ctx* ctx);
/* glGetError-style text of the last failure on this context (ctx may be NULL for srt_create failures). */
const char* srt_last_error(const srt_ctx* ctx);

/* Run on an existing CUDA stream (a cudaStream_t passed as void*); NULL = the context's own stream. */
int srt_set_stream(srt_ctx* ctx, void* cuda_stream);
int srt_set_option(srt_ctx* ctx, int key, int value);

/* AssetUtils::UploadModelDataToGPU  (src/asset_utils/gpu_loader.cpp:135-182): the five SSBOs at
 * binding_offset+0..4, whole-buffer upload; may be called again to replace the scene. */
int srt_upload_scene(srt_ctx* ctx, const srt_bvh* bvhs, uint32_t n_bvhs, const srt_bvh_node* nodes, uint32_t n_nodes,
                     const srt_material* materials, uint32_t n_materials, const srt_triangle* triangles,
                     uint32_t n_triangles, const srt_vertex* vertices, uint32_t n_vertices);

/* AssetUtils::GPUTexture::GPUTexture(file, true) + GetHandle()  (include/asset_utils/gpu_texture.h:24-68,133-137) for the
 * material `material_index` of the uploaded scene: the decoded image, RGBA8, row 0 first, REPEAT / LINEAR like the
 * reference's sampler.  Takes effect with SRT_OPT_TEXTURES = 1; a new srt_upload_scene forgets every texture. */
int srt_set_texture(srt_ctx* ctx, uint32_t material_index, const uint8_t* rgba8, uint32_t width, uint32_t height);

/* AssetUtils::UpdateModelMatrix  (src/asset_utils/gpu_loader.cpp:185-196): glBufferSubData of bvhs[index].frame,
 * 16 floats column-major. */
int srt_update_frame(srt_ctx* ctx, uint32_t index, const float* mat4_column_major);

/* UpdateNoiseTex  (src/main.cpp:281-296): the two RGB32F texture buffers, count texels of 3 floats each.
 * count must equal width*height of the render (the shader indexes them modulo Width*Height). */
int srt_upload_noise(srt_ctx* ctx, const float* unit_vec3, const float* uniform_vec3, uint32_t count);

/* lights SSBO at binding 4 + SetInt("lightCount")  (src/main.cpp:684,688-692).  count <= SRT_MAX_LIGHTS. */
int srt_set_lights(srt_ctx* ctx, const srt_light* lights, int32_t count);

/* Compute::SetBool/SetInt/SetUInt/SetVec3  (src/main.cpp:668-685): the per-frame uniforms.  light_count is
 * taken from srt_set_lights when params->light_count < 0. */
int srt_set_params(srt_ctx* ctx, const srt_params* params);

/* Texture::getTextureHandle / Texture32::getTextureHandle  (src/graphics/texture.cpp:50-97,142-186): allocate
 * the RGBA8 output image and the RGBA32F accumulation image (zero-initialised, decree Q8).
 * width and height must be multiples of 8 (src/main.cpp:706 dispatches W/8 x H/8 groups unguarded). */
int srt_resize(srt_ctx* ctx, int32_t width, int32_t height);

/* Restrict rendering to rows [y0, y1) (multiples of 8) -- the image-tile half of the multi-GPU split.
 * Default is the whole image. */
int srt_set_rows(srt_ctx* ctx, int32_t y0, int32_t y1);
/* Of the 8-row bands inside the row range, render only those with band % world == rank (interleaved bands keep
 * the ranks' loads even when the model covers part of the picture).  Default rank 0 of world 1. */
int srt_set_stripe(srt_ctx* ctx, int32_t rank, int32_t world);

/* glDispatchCompute(W/8,H/8,1) + glMemoryBarrier + glFinish  (src/main.cpp:706-718), n_frames times:
 * renders frames accum_frames .. accum_frames+n_frames-1 (1 sample per pixel each), accumulating in
 * frame order, then advances params.accum_frames by n_frames.  With params.reset_accum set it clears the
 * accumulation image instead (raytrace_compute.glsl:390-393).  Returns after the work completed. */
int srt_render(srt_ctx* ctx, int32_t n_frames);
/* Same, without the trailing synchronisation (work is queued on the context's stream). */
int srt_render_async(srt_ctx* ctx, int32_t n_frames);
int srt_sync(srt_ctx* ctx);

/* Texture::updateImageDataFromGPU / Texture32::updateImageDataFromGPU  (src/graphics/texture.cpp:99-106,192-200). */
int srt_read_output(srt_ctx* ctx, uint8_t* rgba8 /* W*H*4 */);
int srt_read_accum(srt_ctx* ctx, float* rgba32f /* W*H*4 */);
/* Replace the accumulation image (resume from a checkpoint / install a reduced multi-GPU sum). */
int srt_write_accum(srt_ctx* ctx, const float* rgba32f /* W*H*4 */);
/* Re-run only the tonemap (raytrace_compute.glsl:412-413) from the current accumulation image with
 * divisor accum_frames; used after a multi-GPU reduce. */
int srt_tonemap(srt_ctx* ctx, int32_t accum_frames);
/* Device address of the RGBA32F accumulation image (W*H*4 floats) so that a collective (NCCL through
 * torch.distributed) can reduce it in place.  Owned by the context. */
int srt_accum_device_ptr(srt_ctx* ctx, void** device_ptr, uint64_t* n_floats);

/* Parity aid: primary-hit primitive per pixel of the last frame rendered (sphere index or global triangle
 * index, SRT_NO_HIT for a miss).  Recording is switched on by the first call with enable=1. */
int srt_enable_primary_ids(srt_ctx* ctx, int enable);
int srt_read_primary_ids(srt_ctx* ctx, uint32_t* ids /* W*H */);

/* AssetUtils::UpdateRays + the test kernel of ray_intersects.glsl:135-161  (src/asset_utils/gpu_loader.cpp:198-210,
 * include/compute/tests/BVH_intergration_tests.cpp:69-94): closest hit of each ray against the first bvh_count
 * BVHs; hit_tri[i] = global triangle index or SRT_NO_HIT, hit_t[i] = final intersection_distance (may be NULL). */
int srt_trace_rays(srt_ctx* ctx, const srt_ray* rays, uint32_t n_rays, uint32_t bvh_count, uint32_t* hit_tri,
                   float* hit_t);

/* IntersectionUtils::BVH<Prim>::BVH(prims, center_fn, bounds_fn)  (include/intersection_utils/bvh.h:40-148), built on
 * the device: the same midpoint splits, the same unstable partition, the same node numbering -- byte for byte the
 * arrays GetBVH() / GetPrims() of the reference's host builder give.
 * centre_xyz / min_xyz / max_xyz: what center_fn and bounds_fn return per primitive, n_prims xyz triples each (host).
 * nodes_out: room for 2 * n_prims - 1 nodes; children and first primitives are relative to this tree (the flattening
 * of several models rebases them, gpu_loader.cpp:120-130).  order_out[k] = input index of the primitive that ends
 * up k-th (GetPrims()[k]).  depth_out / device_ms_out may be NULL. */
int srt_build_bvh(srt_ctx* ctx, const float* centre_xyz, const float* min_xyz, const float* max_xyz, uint32_t n_prims,
                  srt_bvh_node* nodes_out, uint32_t* n_nodes_out, uint32_t* order_out, uint32_t* depth_out,
                  float* device_ms_out);

/* srt_trace_rays without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, same results (bit for bit) as srt_trace_rays, but the device staging of the rays and of the two result
 * arrays lives in a file-scope `static __device__` array inside the library instead of per-call cudaMalloc'ed buffers:
 *   - at most SRT_TRACE_STATIC_MAX_RAYS rays per call; more is rejected with an error that names the limit;
 *   - the array is 40 B x SRT_TRACE_STATIC_MAX_RAYS = 2.5 MiB and every GPU that loads the library holds it until exit;
 *   - the call makes no cudaMalloc / cudaFree / allocating library call (copies, two memsets, one kernel, a synchronise);
 *   - every call on a device shares the array, which keeps the last call's contents: the result words are cleared before
 *     the kernel runs, and calls whose GPU work OVERLAPS on one device (two contexts or streams) race -- serialise them. */
#define SRT_TRACE_STATIC_MAX_RAYS 65536u
int srt_trace_rays_static(srt_ctx* ctx, const srt_ray* rays, uint32_t n_rays, uint32_t bvh_count, uint32_t* hit_tri,
                          float* hit_t);

/* ---- several B200s behind one srt_render (no reference call site: src/main.cpp:706 is a single dispatch) ---------------
 * the frame shards by image tile and by sample range with full replicas of scene and noise tables on
 * every GPU and no exchange inside the path loop; ONE collective over NVLink per srt_render call.  In a group
 *   - srt_render / srt_render_async render this rank's share and then run the collective + rank 0's tonemap, all on the
 *     context's stream;
 *   - rank 0 holds the picture: srt_read_output / srt_read_accum / srt_accum_device_ptr / srt_tonemap there act on the
 *     GATHERED accumulation image; every rank's own image only holds what that rank rendered, so progressive rendering
 *     without resets never counts a contribution twice;
 *   - every rank must be given the same scene, noise tables, lights, uniforms and image size, and make the same calls.
 * SRT_OPT_SPLIT picks the partition:
 *   SRT_SPLIT_BANDS   (default) rank r renders the interleaved 8-row bands b = r (mod world); the collective gathers every
 *                     rank's own rows (Width*Height*16/world bytes each) on rank 0: the bits of a single-GPU render
 *   SRT_SPLIT_SAMPLES rank r renders the whole image for the r-th of `world` consecutive ranges of the call's frames; the
 *                     collective is one ncclReduce(sum) of the fp32 images: fp32 additions re-ordered, equal to the
 *                     single-GPU image within tolerance (1e-4 relative), not in bits
 * NCCL is loaded with dlopen("libnccl.so.2") when the first group is formed; without it these calls fail, the rest of
 * the library does not need it. */
#define SRT_OPT_SORT_RAYS 16       /* wavefront schedule: 1 = before a wave is traced its two queues are regrouped by the cell the
                                     rays start in (15-bit Morton code in the first BVH's root box; counting sort, 8 small
                                     launches).  Another order of answering rays, never another result (parity-tested).
                                     Default 0: measured on configs[4] (10 M-triangle soup) it is SLOWER, 354 vs 215 ms per
                                     step -- the queues come out of the logic kernel in pixel order, which is already a finer
                                     grouping by origin than any cell grid; what diverges there is the rays' directions */
#define SRT_OPT_DEVICE_RELAYOUT 15 /* where srt_upload_scene turns the reference's arrays into this library's records: 0 = a serial walk
                                     on the host, 1 = kernels on the device (the raw arrays are copied over as they are: DMA
                                     when the caller's arrays are page-locked), 2 (default) = on the device for scenes of more
                                     than 2 Mi triangles, and for scenes whose hit finder is built on the device (SRT_OPT_DEVICE_FINDER_BUILD,
                                     16 Ki triangles and more: the finder build goes on from the re-layout's sweeps, nothing comes back).  The device version needs the reference's node numbering (children
                                     after their parent, bvh.h:96-131) and valid indices; anything else goes to the host walk,
                                     which handles or reports it.  Same records up to where a node pair sits in its array: no
                                     ray result depends on it (parity-tested).  srt_get_upload_info says what the last upload did */
#define SRT_OPT_DEVICE_FINDER_BUILD 17 /* 1 (default) = the hit finder's private trees (SRT_OPT_FINDER) are built by kernels at srt_upload_scene
                                     time, a tree level per turn (csrc/srt_finder_build.cu), for scenes of 16 Ki triangles and more; 0 = always on host threads.  Either tree gives
                                     the same pictures: the finder only collects candidates, the replay of the shader's own culls
                                     decides (parity-tested both ways) */
#define SRT_OPT_OVERLAP 18          /* wavefront schedule: 1 (default) = two sets of path slots; a chunk's waves run on an internal stream and
                                     its tail kernel + resolve on the context's stream, so the tail of one chunk -- or of one srt_render
                                     call -- runs beside the waves of the next.  The accumulation image and the output are only ever
                                     touched on the context's stream, in call order: same bits.  0 = everything on the context's stream,
                                     one chunk after the other.  Off by itself while primary ids are recorded or tests are counted */
#define SRT_OPT_FUSED_PRIMARY 19    /* wavefront schedule with the hit finder: 1 = no separate init kernel where every path of a launch starts
                                     in its first wave; that wave's find kernel makes the camera rays itself and only a ray that comes back
                                     with candidate triangles becomes a path in a slot -- any other ends there with the sky sample, one bit
                                     instead of slot records and a trip through the logic kernel.  Same bits (parity-tested both ways).
                                     0 (default) = init kernel, then the waves: measured FASTER (Rubik 1080p x 64 spp 26.3 against 29.3 ms,
                                     an eighth of the frame 5.24 against 5.43): the two thirds of the camera rays that miss the root box
                                     cost the persistent find kernel 2.5x what they cost the init kernel (half the occupancy, its fetch /
                                     flush loop per 32 paths), more than the saved records give back.  Kept as the measured alternative.
                                     Off by itself while tests are counted or rays sorted */
#define SRT_OPT_SPLIT 14
#define SRT_SPLIT_BANDS 0
#define SRT_SPLIT_SAMPLES 1
#define SRT_GROUP_ID_BYTES 128
/* One process per GPU (torchrun / mpirun): rank 0 calls srt_group_unique_id, the launcher's own channel carries the
 * SRT_GROUP_ID_BYTES bytes to every rank, every rank calls srt_group_join with them (collective: returns when all have). */
int srt_group_unique_id(void* id, uint64_t bytes);
int srt_group_join(srt_ctx* ctx, int32_t rank, int32_t world, const void* id, uint64_t bytes);
/* One process, n GPUs: *out is a HEAD context (rank 0, devices[0]).  Options, uploads, uniforms, resize, render, sync
 * and counter calls on it act on every device (srt_render on one host thread per device); reads come from rank 0;
 * srt_destroy(head) destroys all.  n = 1 is srt_create. */
int srt_create_multi(srt_ctx** out, const int* devices, int32_t n);
/* Sum of the CUDA-event durations of this rank's collectives since the last call, and how many there were. */
int srt_get_collective_ms(srt_ctx* ctx, double* ms, uint64_t* calls);

/* What the last srt_upload_scene did: *on_device = 1 when the records were built by kernels (SRT_OPT_DEVICE_RELAYOUT),
 * *device_ms = their CUDA-event time.  No reference counterpart (the reference uploads its arrays as they are). */
int srt_get_upload_info(srt_ctx* ctx, int32_t* on_device, float* device_ms);

int srt_get_counters(srt_ctx* ctx, srt_counters* out);
/* Sums of the CUDA-event durations of the wavefront schedule's trace and logic launches since the last call
 * (needs SRT_OPT_TIME_KERNELS = 1); waves = launches of each kind.  Synchronises the stream. */
int srt_get_kernel_times(srt_ctx* ctx, double* trace_ms, double* logic_ms, uint64_t* waves);
int srt_reset_counters(srt_ctx* ctx);
/* Hit finder (SRT_OPT_FINDER): whether the next srt_render of the uploaded scene goes through it, and the two figures
 * srt_upload_scene measured on its sample rays (box tests + 1.5 x triangle tests per ray, walking the reference BVH
 * vs finder + replay).  No reference call site (the shader has one way to intersect).  Any pointer may be NULL. */
int srt_get_finder_info(srt_ctx* ctx, int32_t* active, double* cost_reference, double* cost_finder);

/* Measurement aid (no reference call site): bytes per second this GPU delivers for RANDOM 64-byte records -- the size of a
 * device BVH node pair -- gathered from an array of `bytes`, with the trace kernel's launch shape.  The array size
 * picks the level measured: <= 64 KiB stays in every SM's L1 (Rubik's 70 KB scene), 16-64 MiB in the 126 MB L2
 * (the stand-in's 8 MB scene, the noise tables), far above 126 MB is HBM (configs[4]).  16 KiB .. 64 GiB.
 * chase = 1: each address depends on the record loaded before (a traversal's node -> child chain);
 * chase = 0: independent addresses.  These are the roofline denominators bench.py quotes. */
int srt_measure_gather(srt_ctx* ctx, uint64_t bytes, int32_t chase, double* gbytes_per_s);

/* Measurement aid (no reference call site): thread-instructions per second of UN-FUSED fp32 arithmetic on this GPU
 * (the library is built -fmad=false: the numeric decree allows only explicit fma, so box and triangle tests are FADD /
 * FMUL / FMNMX, not FFMA).  mix 0 = FMUL + FADD, mix 1 = FMUL + FADD + 2 FMNMX (the slab test's proportions).  One
 * fp32 operation per thread-instruction: the result is also the FP32 op/s roof for such code.  Nominal: 148 SMs x 4
 * sub-partitions x 32 lanes x clock. */
int srt_measure_issue(srt_ctx* ctx, int32_t mix, double* thread_ginst_per_s);

/* Checked build (libsrt_b200_checked.so, built -DSRT_CHECKED: the wavefront kernels bounds-check every stack push and
 * every queue / slot / sample index into srt_counters.check_flags; compute-sanitizer is not available on the pool this
 * was developed on).  srt_check_selftest raises bit 31 of check_flags through the same macro, which proves the checks are
 * compiled in; in the normal build it launches an empty kernel and check_flags stays 0.  No reference call site. */
int srt_check_selftest(srt_ctx* ctx);

/* Device-side evaluation of the numeric decree (srt_detmath.h), for bit-for-bit checks against the host:
 * which: 0 = sin, 1 = cos, 2 = pow(x, 1/2.4), 3 = randFloat(x, y[i]) (y required), 4 = linearToSrgb. */
int srt_eval_detmath(srt_ctx* ctx, int which, const float* x, const float* y, float* out, uint32_t n);

/* Library/build identification, e.g. "srt_b200 0.1 sm_100a". */
const char* srt_version(void);

#ifdef __cplusplus
}
#endif

#endif /* SRT_B200_H_ */
