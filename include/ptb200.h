/*
 * ptb200.h — C ABI of libptb200.so, the B200 (sm_100a) CUDA backend that stands where the
 * reference's Vulkan compute pipeline stood.
 *
 * The reference has no FFI of its own for this path: the seam is the compute-pipeline
 * contract inside RayTracing/Graphics/Renderer.cs.  Each entry point below
 * cites the Renderer member (file:line under /root/reference/RayTracing/) it replaces.
 * The C# binding a maintainer would add is shown in INTEGRATION.md and shipped as source in
 * pathtracing_b200/cs/NativeTracer.cs.
 *
 * Conventions
 *   - plain pointers and sizes only; every struct is blittable (P/Invoke / ctypes friendly);
 *   - every call returns an int status: 0 = PTB_OK, negative = ptb_status; nothing throws or
 *     aborts across the ABI (the reference throws `new Exception(string)` on any non-Success
 *     Vk result, e.g. Renderer.cs:115-118; the C# wrapper rethrows from ptb_last_error);
 *   - a context is externally synchronised (one host thread at a time, like the Vulkan
 *     objects it replaces); distinct contexts may be used from distinct threads;
 *   - launches are asynchronous on the context's stream; ptb_wait / ptb_read_* synchronise;
 *   - there is NO CPU fallback: without a CUDA device ptb_create fails with PTB_ERR_CUDA.
 */
#ifndef PTB200_H
#define PTB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PTB_ABI_VERSION 1

typedef enum ptb_status {
    PTB_OK = 0,
    PTB_ERR_INVALID = -1,     /* bad argument (NULL, zero size, bad flags, pitch too small) */
    PTB_ERR_CUDA = -2,        /* CUDA runtime error; text in ptb_last_error */
    PTB_ERR_NOMEM = -3,       /* device or host allocation failed */
    PTB_ERR_STATE = -4,       /* call not valid in the current state (e.g. resolve with 0 spp) */
    PTB_ERR_UNSUPPORTED = -5  /* feature not built / size beyond limits */
} ptb_status;

typedef struct ptb_context ptb_context;   /* opaque; owns all device memory and its stream */

/* ptb_config.flags */
#define PTB_CFG_KEEP_F32      0x1u  /* ptb_compute_frame also stores the float4 image (Test.hlsl:1 RWTexture2D<float4>) */
#define PTB_CFG_KEEP_HIT_IDS  0x2u  /* ptb_compute_frame also stores int32 hit ids (0 = sphere, -1 = miss) */
#define PTB_CFG_EXTERNAL_STREAM 0x4u /* use `stream` (a cudaStream_t) instead of creating one */
#define PTB_CFG_FORCE_IEEE    0x8u  /* always run the literal IEEE-intrinsic kernels (never the shortened, proven-equal
                                       FMA sequences); results are bit-identical either way — for tests / A-B timing */

#define PTB_CFG_SPIRV_ARITH   0x10u /* evaluate the operation list of the reference's CHECKED-IN compiled module
                                       Assets/Shaders/Compiled/Ray/Test.spirv (what Renderer.cs:356-357 loads) instead of
                                       the source order of Test.hlsl: uv by the folded reciprocal fl(1/1080) [general:
                                       fl(1/div)], disc = Fma(b, b, a*-3), -b-s = Fma(d.z, -2, -s).  Literal scene only
                                       (camera / sphere are folded into the module).  Same hit mask as the default
                                       arithmetic, radiance within 3.4e-5 relative at 1080p; bit-identical to the numpy
                                       evaluation of the module's words (tests/test_spirv_pin.py, test_gpu_spirv.py) */

/* Parameters that are literals in the reference (Test.hlsl:6,8,12,13; App.cs:27).
 * ptb_default_config fills them so that width=1920,height=1080 reproduces the reference. */
typedef struct ptb_config {
    uint32_t struct_size;       /* = sizeof(ptb_config); checked */
    int32_t  device;            /* CUDA ordinal */
    uint32_t width, height;     /* image size (Renderer.cs:128-129 Window.FramebufferSize) */
    float    uv_div_x, uv_div_y;/* Test.hlsl:6 `resolution`; default (height,height) */
    float    cam[3];            /* Test.hlsl:8 */
    float    centre[3];         /* Test.hlsl:12 */
    float    radius;            /* Test.hlsl:13 */
    uint32_t seed;              /* RNG seed for jittered samples (no reference counterpart) */
    uint32_t max_depth;         /* bounce limit for scene rendering; reference = 0 */
    uint32_t flags;             /* PTB_CFG_* */
    void    *stream;            /* cudaStream_t when PTB_CFG_EXTERNAL_STREAM */
} ptb_config;

/* Fill *cfg with the reference literals for a width x height image on device 0. */
int ptb_default_config(ptb_config *cfg, uint32_t width, uint32_t height);

/* Library / ABI version (PTB_ABI_VERSION of the built library). */
int ptb_abi_version(void);

/* Number of CUDA devices visible, or a negative ptb_status. */
int ptb_device_count(void);

/* ---- lifecycle -------------------------------------------------------------------------
 * ptb_create  replaces Renderer.CreateResources (Renderer.cs:105-196: the W x H RGBA8 storage
 *             image, its memory/view, layout transition) + Renderer.CreateComputePipeline
 *             (Renderer.cs:293-403: descriptor layout, SPIR-V module, pipeline, ComputeFence).
 * ptb_destroy replaces the compute part of Renderer.Dispose (Renderer.cs:1192-1216), and —
 *             unlike the reference, which leaks its compute objects — frees everything.
 *             ptb_destroy(NULL) is a no-op returning PTB_OK. */
int ptb_create(const ptb_config *cfg, ptb_context **out_ctx);
int ptb_destroy(ptb_context *ctx);

/* UTF-8 text of the last error on this context (ctx == NULL: last ptb_create /
 * argument error on the calling thread).  Never NULL; "" when no error. */
const char *ptb_last_error(const ptb_context *ctx);

/* ---- the per-frame hot path ------------------------------------------------------------
 * ptb_compute_frame replaces Renderer.ComputeFrame (Renderer.cs:1006-1040): one launch of the
 *             CSMain-equivalent kernel (Test.hlsl:3-40) over the whole image with an explicit
 *             bounds guard (the reference over-dispatches 1920x1088 and relies on the API
 *             dropping OOB writes).  Output: the context's RGBA8 UNORM image (the effective
 *             store format, Renderer.cs:124), plus float4 / hit ids when configured.
 * ptb_wait    replaces WaitForFences(ComputeFence, inf) (Renderer.cs:972). */
int ptb_compute_frame(ptb_context *ctx);
int ptb_wait(ptb_context *ctx);
/* n_frames consecutive ptb_compute_frame's as ONE persistent launch (literal scene, RGBA8 output only, width % 4 == 0, no
 * row shard; anything else falls back to n_frames launches).  The reference's render loop (Renderer.cs:933-1004) calls
 * ComputeFrame once per displayed frame and each frame is ~2 us of GPU work, so a launch per frame is bound by launch
 * latency; here one resident wave of blocks walks the row chunks of every frame back to back.  Every frame is computed
 * and stored in full; all frames write the same image, exactly as the reference's frames overwrite TestImage. */
int ptb_compute_frames(ptb_context *ctx, uint32_t n_frames);

/* Headless consumers of the image (replace the fragment-stage sampling of ComputeImage,
 * Renderer.cs:1086-1089).  pitch_bytes = host row stride; 0 means tightly packed.
 * These synchronise the context's stream. */
int ptb_read_rgba8(ptb_context *ctx, void *host_dst, size_t pitch_bytes);
/* Non-blocking variant: the copy is queued behind the work submitted so far on a separate copy
 * stream, so the next frame's kernels overlap the PCIe transfer (the reference overlaps compute
 * with presentation the same way: the graphics queue consumes frame k while the compute queue may
 * start k+1).  host_dst should be page-locked; it is valid after the next ptb_wait.  The next
 * writer of the image (ptb_compute_frame / ptb_resolve) is ordered after the copy. */
int ptb_read_rgba8_async(ptb_context *ctx, void *host_dst, size_t pitch_bytes);
int ptb_read_f32(ptb_context *ctx, float *host_dst, size_t pitch_bytes);      /* needs KEEP_F32 */
int ptb_read_hit_ids(ptb_context *ctx, int32_t *host_dst, size_t pitch_bytes);/* needs KEEP_HIT_IDS */

/* Zero-copy access for device-side consumers (CUDA device pointers, valid until destroy). */
void *ptb_device_rgba8(ptb_context *ctx);
void *ptb_device_f32(ptb_context *ctx);
void *ptb_device_accum(ptb_context *ctx);
void *ptb_device_peer_flags(ptb_context *ctx);   /* this context's rendezvous flag block (ptb_peer_barrier) */

/* ---- multi-sample accumulation (north_star stages n1/n2; no reference counterpart) ------
 * ptb_render_samples: for each pixel, samples s in [first_sample, first_sample+n_samples)
 *   are evaluated in-kernel with sub-pixel jitter from the counter-based RNG keyed
 *   (pixel, s, seed), summed sequentially, and the partial is added to the float4
 *   accumulation buffer.  jitter == 0 evaluates the literal CSMain ray n_samples times.
 * ptb_resolve: image = UNORM8(accum / spp_total) (flags: PTB_RESOLVE_*), written to the
 *   context's RGBA8 image. */
#define PTB_RESOLVE_SRGB 0x1u   /* apply the sRGB OETF (what a B8G8R8A8Srgb target does, SwapChain.cs:158) */
#define PTB_RESOLVE_BGRA 0x2u   /* B,G,R,A byte order */
int ptb_accum_reset(ptb_context *ctx);
int ptb_render_samples(ptb_context *ctx, uint32_t first_sample, uint32_t n_samples, int jitter);
int ptb_resolve(ptb_context *ctx, uint32_t flags);
/* ptb_render_samples followed by ptb_resolve, as ONE kernel launch where that is possible (literal scene, no row shard,
 * no PTB_RESOLVE_SRGB): the accumulate kernel resolves each pixel while its sum is still in registers, which saves the
 * resolve pass's read of the accumulation buffer and a launch per frame.  Same results as the two calls, bit for bit.
 * The fused form writes a SECOND image buffer and makes it the current one, so that an asynchronous read-back of the
 * previous frame (ptb_read_rgba8_async) keeps overlapping the next frame's rendering: re-query ptb_device_rgba8 after
 * every call, and do not combine it with an image exported through ptb_ipc_export. */
int ptb_render_resolve(ptb_context *ctx, uint32_t first_sample, uint32_t n_samples, int jitter, uint32_t flags);
uint32_t ptb_accum_spp(const ptb_context *ctx);
/* Checkpoint / restore of the progressive state (float4 accumulation + spp). */
int ptb_read_accum(ptb_context *ctx, float *host_dst, size_t pitch_bytes);
int ptb_write_accum(ptb_context *ctx, const float *host_src, size_t pitch_bytes, uint32_t spp_total);
/* Use caller-owned device memory (W*H*4 floats, 16 B aligned) as the accumulation buffer,
 * e.g. a buffer registered with a collective library.  NULL restores the internal one. */
int ptb_set_accum_buffer(ptb_context *ctx, void *device_ptr);
/* Multi-GPU image-tile sharding (north_star stage n8, partitioning A): this context renders only
 * rows  row_offset + i*row_stride  (interleaved for load balance: hits cluster in the image).
 * Untouched rows of the image / accumulation buffer keep their contents (zero after create /
 * ptb_accum_reset), so the per-device results merge by a sum == gather of disjoint rows and the
 * merged image is bit-identical to a single-device render.  Default (0,1) = every row. */
int ptb_set_row_shard(ptb_context *ctx, uint32_t row_offset, uint32_t row_stride);
/* The same with bands of rows_per_band consecutive rows as the interleaving unit: this context renders the bands
 * g = row_offset + i*row_stride, band g = rows [g*rows_per_band, (g+1)*rows_per_band) clipped to the image.
 * The scene path traces primary rays in 8x4 pixel tiles, so bands of a multiple of 4 rows keep a warp's rays
 * adjacent (single interleaved rows spread a tile over 4*row_stride image rows). rows_per_band = 1 == ptb_set_row_shard. */
int ptb_set_row_shard_bands(ptb_context *ctx, uint32_t row_offset, uint32_t row_stride, uint32_t rows_per_band);
/* Override the spp count used by ptb_resolve after an external merge of accum buffers. */
int ptb_set_accum_spp(ptb_context *ctx, uint32_t spp_total);

/* ---- scene path (north_star stages n3-n7; NO reference counterpart: the reference's scene is the
 * literals of Test.hlsl:8,12-13 and its only intersection is Test.hlsl:16-27) -------------------
 * ptb_upload_scene copies a scene description from HOST memory, packs it into 16 B-aligned device
 * arrays (spheres: float4 centre+radius; triangles: 3 x float4 = v0|id, e1|material, e2|0, stored
 * in BVH leaf order), and builds the acceleration structures ON THE GPU: Morton-code LBVH (63-bit
 * codes, radix sort, Karras hierarchy), bottom-up AABB refit, SAH leaf collapse, flattened into
 * 64 B two-child nodes.  One hierarchy per primitive class (spheres / triangles).
 * While a scene is resident, ptb_compute_frame renders sample 0 of every pixel without jitter
 * (RGBA8 + optional float4 + primary-ray primitive ids: spheres 0..ns-1, triangles ns..ns+nt-1,
 * -1 = miss) and ptb_render_samples runs the wavefront path tracer (one wave = up to 64 samples of
 * every pixel of the row shard, see ptb_scene_set_wave_samples; per bounce one fused closest-hit +
 * scatter kernel; live paths are repacked with warp ballot + prefix sum).  ptb_clear_scene returns the context to the literal CSMain path.
 * The arithmetic spec (ray generation = Test.hlsl:7-10 generalised to a camera basis, sphere
 * quadratic = Test.hlsl:16-27 + far root, Moeller-Trumbore triangles, Lambert / metal /
 * dielectric scattering, RNG keys) is frozen in oracle/pt_scene_oracle.h. */
#define PTB_MAT_LAMBERT    0u
#define PTB_MAT_METAL      1u
#define PTB_MAT_DIELECTRIC 2u
#define PTB_SHADE_PATH     0u   /* multi-bounce paths lit by the sky gradient */
#define PTB_SHADE_NORMALS  1u   /* Test.hlsl:28-36: hit -> n*0.5+0.5, miss -> (uv.x, uv.y, 0) */

typedef struct ptb_material {   /* 32 bytes */
    uint32_t type;              /* PTB_MAT_* */
    float    albedo[3];
    float    fuzz;              /* metal */
    float    ior;               /* dielectric */
    float    pad[2];
} ptb_material;

typedef struct ptb_scene_desc {
    uint32_t struct_size;       /* = sizeof(ptb_scene_desc); checked */
    uint32_t n_spheres;
    const float    *spheres;        /* host, n_spheres x (cx,cy,cz,r) */
    const uint32_t *sphere_material;/* host, n_spheres */
    uint32_t n_triangles;
    uint32_t n_materials;
    const float    *tri_vertices;   /* host, n_triangles x (v0,v1,v2) = 9 floats */
    const uint32_t *tri_material;   /* host, n_triangles */
    const ptb_material *materials;  /* host, n_materials (>= 1) */
    float    cam_origin[3];         /* Test.hlsl:8 */
    float    cam_u[3], cam_v[3], cam_w[3]; /* ray = normalize(uv.x*U + uv.y*V + W); Test.hlsl:10 = (1,0,0),(0,1,0),(0,0,-1) */
    float    sky_horizon[3], sky_zenith[3];/* escaped paths: lerp by (d.y+1)/2 */
    float    t_min;                 /* ray epsilon, e.g. 1e-3 */
    uint32_t shade_mode;            /* PTB_SHADE_* */
    uint32_t max_leaf;              /* SAH leaf collapse limit (1..8); 0 = default 4 */
    uint32_t flags;                 /* PTB_SCENE_* */
} ptb_scene_desc;
#define PTB_SCENE_COUNT_STATS 0x1u  /* run the counting build of the kernels (ptb_scene_stats) */

typedef struct ptb_scene_info {
    uint32_t n_spheres, n_triangles;
    uint32_t sphere_nodes, triangle_nodes;   /* 64 B nodes after leaf collapse */
    uint32_t sphere_leaves, triangle_leaves;
    uint32_t max_leaf;
    float    build_ms;                       /* GPU time of the whole build (CUDA events) */
    uint64_t device_bytes;                   /* nodes + primitives + materials */
} ptb_scene_info;

/* Work counters of the counting build, summed over every ray since upload / ptb_scene_stats_reset. */
typedef struct ptb_scene_stats {
    uint64_t rays;            /* closest-hit queries (path segments) */
    uint64_t nodes_visited;   /* 64 B node fetches */
    uint64_t prims_tested;    /* primitive intersection tests */
    uint64_t scatters;        /* material evaluations */
    uint64_t paths;           /* camera rays started */
    uint64_t max_stack;       /* deepest traversal stack seen */
} ptb_scene_stats;

int ptb_upload_scene(ptb_context *ctx, const ptb_scene_desc *desc);
int ptb_clear_scene(ptb_context *ctx);
int ptb_scene_info_get(ptb_context *ctx, ptb_scene_info *out);
int ptb_scene_stats_get(ptb_context *ctx, ptb_scene_stats *out);
int ptb_scene_stats_reset(ptb_context *ctx);
/* ptb_render_samples on a scene traces WAVES: one wave = `samples_per_wave` samples of every pixel of the row shard
 * carried through all bounces.  0 (default) = as many as fit the path queues (sized for 64 Mi paths, at most 64 samples),
 * so that small frames and 1/N row shards still fill the GPU at deep bounces; n >= 1 caps it.  Scheduling only: the
 * accumulation buffer receives the samples' radiance in sample order whatever the wave size — results do not change
 * by a bit (tests/test_gpu_scene.py). */
int ptb_scene_set_wave_samples(ptb_context *ctx, uint32_t samples_per_wave);
/* Next-event estimation towards ONE directional light ("sun"; no reference counterpart — Test.hlsl has no light): at every
 * hit that scatters on a PTB_MAT_LAMBERT material the path tracer casts a shadow ray (ANY-HIT traversal of both
 * hierarchies, (t_min, +inf)) along `dir_to_light` (used as given; pass a unit vector) and, when nothing occludes it
 * and c = n.dir > 0 for the normal turned against the incoming ray, adds (throughput*albedo)*(radiance*c) to the
 * path's radiance.  The sky does not contain the sun, so nothing is counted twice.  radiance = (0,0,0) switches it off
 * (the state after ptb_upload_scene).  Spec: oracle/pt_scene_oracle.h "sun"; results are bit-identical to it. */
int ptb_scene_set_sun(ptb_context *ctx, const float dir_to_light[3], const float radiance[3]);
/* Live paths entering each bounce of the LAST wave rendered (out[0] = camera rays, out[d] = paths that survived d
 * scatters); writes min(cap, max_depth + 1) entries and their number to *n.  Synchronises. */
int ptb_scene_queue_sizes(ptb_context *ctx, uint32_t *out, uint32_t cap, uint32_t *n);
/* Download one flattened hierarchy for inspection: which = 0 spheres, 1 triangles.
 * nodes_out: cap_nodes x 16 floats (64 B nodes, layout in csrc/scene.cuh), order_out: cap_prims
 * primitive ids in leaf order.  Either may be NULL; counts are always written. */
int ptb_scene_read_bvh(ptb_context *ctx, int which, void *nodes_out, uint32_t cap_nodes, uint32_t *n_nodes,
                       uint32_t *order_out, uint32_t cap_prims, uint32_t *n_prims);
/* Ray queries against the resident scene with HOST ray arrays (n x 3 floats each).
 * mode 0 = closest hit (out_prim = id or -1, out_t = distance or 0), 1 = any hit in (t_min,t_max)
 * (out_prim = 1/0, out_t may be NULL). */
int ptb_trace_rays(ptb_context *ctx, uint64_t n_rays, const float *origins, const float *dirs, float t_min,
                   float t_max, int mode, int32_t *out_prim, float *out_t);
/* ptb_trace_rays for callers that may not allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, results and errors as ptb_trace_rays, but the device staging of the ray and result arrays
 * (32 bytes per ray) lives in a static device array of the library (PTB_TRACE_RAYS_STATIC_MAX x 32 B = 2 MiB, held by
 * every GPU that loads the library until the process exits) instead of four cudaMalloc'ed buffers: the call makes no
 * cudaMalloc / cudaFree and calls no allocating library routine.  n_rays > PTB_TRACE_RAYS_STATIC_MAX is
 * PTB_ERR_UNSUPPORTED (nothing is launched; split the batch or use ptb_trace_rays).
 * All calls on one GPU share the array, which keeps the previous call's contents: the result part is cleared on
 * the stream before each launch, and the call returns only after its results have been copied out, but two calls
 * that overlap on one GPU (two contexts of that device driven from two host threads) race — serialise them. */
#define PTB_TRACE_RAYS_STATIC_MAX 65536u
int ptb_trace_rays_static(ptb_context *ctx, uint64_t n_rays, const float *origins, const float *dirs, float t_min,
                          float t_max, int mode, int32_t *out_prim, float *out_t);

/* ---- fused multi-GPU merge over NVLink peer memory (north_star stage n8; no reference counterpart) -------------
 * With sample-range sharding every GPU holds a partial accumulation buffer of the whole frame.  ptb_peer_merge, called
 * on every GPU for its own slab of pixels [slab_first, slab_first+slab_count), loads the partials of those pixels from
 * all peers (device pointers into peer memory, rank order), adds them in that order, adds the sum to this context's
 * slab of the running total (the partials are NOT cleared: every GPU calls ptb_accum_clear on its own buffer once all
 * merges have finished), resolves (total / spp_total -> UNORM8, PTB_RESOLVE_*) and stores the
 * pixels into dst_image (the RGBA8 image of the destination GPU, also peer memory).  One kernel, no staging buffer.
 * The caller provides the cross-GPU ordering: every peer's samples are rendered before, and nobody renders again until
 * every GPU's merge has finished (ptb_peer_barrier before and after; ptb_peer_frame does all of it).
 * ptb_ipc_export / ptb_ipc_open move the pointers between processes (cudaIpcMemHandle_t, 64 bytes);
 * which = 0 this context's accumulation buffer, 1 its RGBA8 image, 2 its rendezvous flag block (3, 4: the second partial
 * buffer / image of a peer group).  Mappings are closed by ptb_destroy. */
#define PTB_MERGE_BACKGROUND 0x10u  /* ptb_peer_merge flag: small grid (one block per SM) for a merge that overlaps a render */
#define PTB_MAX_PEERS 16
#define PTB_IPC_HANDLE_BYTES 64
int ptb_ipc_export(ptb_context *ctx, int which, void *handle_out);
int ptb_ipc_open(ptb_context *ctx, const void *handle, void **device_ptr_out);
/* Rendezvous of the GPUs taking part in a peer merge, on the context's stream, without the host: a one-warp kernel
 * stores `epoch` into this GPU's slot of every peer's flag block (ptb_ipc_export which = 2; peer_flags in rank order,
 * own block included) and waits until every peer has stored `epoch` into this GPU's block.  phase 0 / 1 are two
 * independent sets of slots (e.g. "rendered" and "merged").  Epochs must increase by one per use of a phase.
 * The flag store is a system-scope release, the wait a system-scope acquire.  Gives up after ~10 s
 * (PTB200_BARRIER_TIMEOUT_MS): ptb_peer_merge / ptb_accum_clear queued behind it then do nothing and the next ptb_wait
 * returns PTB_ERR_STATE. */
int ptb_peer_barrier(ptb_context *ctx, uint32_t n_peers, void *const *peer_flags, uint32_t self_rank, uint32_t phase,
                     uint32_t epoch);
int ptb_accum_clear(ptb_context *ctx);   /* zero the accumulation buffer only (keeps the slab total and the spp count);
                                            a no-op on the device when a rendezvous before it timed out */
int ptb_peer_merge(ptb_context *ctx, uint32_t n_peers, void *const *peer_accum, void *dst_image, uint32_t slab_first,
                   uint32_t slab_count, uint32_t spp_total, uint32_t flags);

/* ---- native multi-GPU frame loop: the calls above composed inside the library (one call per frame and rank) ----------
 * Sample-range sharding with the fused merge, pipelined: frame k renders into partial buffer k & 1 on the context's stream
 * while a second, highest-priority stream merges frame k-1 (rendezvous, ptb_peer_merge into image k & 1 of the
 * destination rank, rendezvous, clear of the partial).  No host synchronisation and no NCCL on the data path.
 *   ptb_peer_group_create   allocates the second partial buffer, the second image, the flag block and the merge stream;
 *   ptb_ipc_export          which = 0 partial[0], 1 image[0], 2 flag block, 3 partial[1], 4 image[1] (after create);
 *   ptb_peer_group_local    the same five pointers for peers living in the SAME process (CUDA IPC cannot map a
 *                           process's own allocations);
 *   ptb_peer_group_bind     rank-ordered pointer tables (own entries included) + the destination rank's two images;
 *                           this rank merges the pixel slab [rank*P/n ...) of every frame (split as dist.py sample_range).
 *                           Restarts the rendezvous epochs: after every rank has bound, synchronise the ranks once on
 *                           the host before the first ptb_peer_frame (also the way to recover from a timed-out peer);
 *   ptb_peer_frame          renders THIS rank's samples [first_sample, first_sample + n_samples) of the frame (literal or
 *                           uploaded scene, row shards honoured) and queues its merge; spp_total = samples per pixel in
 *                           the merged image after this frame, over all ranks; flags = PTB_RESOLVE_*.  On the
 *                           destination rank the merged frame becomes the current image: ptb_read_rgba8_async delivers
 *                           it once its merge has finished, overlapping the next frame's rendering;
 *   ptb_peer_finish         makes the context's stream wait for every merge queued so far;
 *   ptb_peer_read_total     this rank's slab of the merged float4 accumulation (host copy; parity checks).
 * A rendezvous that times out (PTB200_BARRIER_TIMEOUT_MS, default 10 s) raises a device-side error word: the merge and
 * clear kernels behind it do nothing, ptb_wait returns PTB_ERR_STATE, ptb_accum_reset clears the condition. */
#define PTB_PEER_RENDER_ONLY 0x20u   /* ptb_peer_frame flag: render into the frame's partial buffer, do not merge yet (a frame
                                        built from several sample ranges; the last call, possibly with n_samples = 0, merges) */
int ptb_peer_group_create(ptb_context *ctx);
void *ptb_peer_group_local(ptb_context *ctx, int which);
int ptb_peer_group_bind(ptb_context *ctx, uint32_t n_peers, uint32_t self_rank, uint32_t dst_rank, void *const *partial0,
                        void *const *partial1, void *const *flags, void *dst_image0, void *dst_image1);
int ptb_peer_frame(ptb_context *ctx, uint32_t first_sample, uint32_t n_samples, int jitter, uint32_t spp_total, uint32_t flags);
int ptb_peer_finish(ptb_context *ctx);
int ptb_peer_read_total(ptb_context *ctx, float *host_dst, uint32_t *slab_first, uint32_t *slab_count);

/* ---- present-equivalent -----------------------------------------------
 * ptb_present replaces the full-screen-quad draw (Renderer.cs:1079-1091) + PSMain
 *   (PixelShader.hlsl:8-19): nearest/repeat sampling of the RGBA8 image onto a dst_w x dst_h
 *   BGRA8 target (SwapChain.cs:154-165), optional sRGB OETF (PTB_RESOLVE_SRGB).  Headless. */
int ptb_present(ptb_context *ctx, uint32_t dst_w, uint32_t dst_h, uint32_t flags);
int ptb_read_present(ptb_context *ctx, void *host_dst, size_t pitch_bytes);

/* ---- measurement -----------------------------------------------------------------------
 * CUDA-event timing on the context's own stream (torch.cuda.Event only sees torch's stream). */
int ptb_timer_begin(ptb_context *ctx);
int ptb_timer_end(ptb_context *ctx, float *elapsed_ms);     /* synchronises */
/* Kernels launched by this context since creation (for bench.py's gpu_launches). */
uint64_t ptb_launch_count(const ptb_context *ctx);
/* Which CSMain kernel variant this context runs: 0 = literal IEEE intrinsics, 1 = shortened exact
 * sequences, 2 = shortened + camera-on-axis specialisation — all three produce identical bits; 3 = the as-compiled
 * operation list of Test.spirv (PTB_CFG_SPIRV_ARITH). */
int ptb_kernel_mode(const ptb_context *ctx);
/* Proof by exhaustion, on the device, that a shortened sequence equals the IEEE intrinsic on its
 * whole domain: which = 0 sqrt over [2^-20,2^24), 1 reciprocal over [2^-12,2^12), 2 division of
 * every float in [2^-24,2^17) by `divisor` (5 = the same for the 2-instruction split-reciprocal division), 3 = the per-pixel "every sample of this pixel misses" classification of THIS
 * context's parameters against the exact discriminant at an 8x8 grid of jitter positions (both extremes included) of
 * every pixel so classified; 4 = the per-row span table of the 1 spp frame kernels (quads of four pixels that skip the
 * intersection): every pixel of every skipped quad against the exact discriminant.  Writes the number of disagreeing inputs (must be 0
 * for the fast kernels to be selected) and the number of inputs tested. */
int ptb_selftest_exact_math(ptb_context *ctx, int which, float divisor, uint64_t *mismatches, uint64_t *tested);

#ifdef __cplusplus
}
#endif
#endif /* PTB200_H */
