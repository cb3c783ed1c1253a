// scene.cu — C-ABI glue of the scene path (include/ptb200.h: ptb_upload_scene ... ptb_trace_rays)
// and its hooks into ptb_compute_frame / ptb_render_samples (north_star stages n3-n7; the
// reference's "scene" is the literals of RayTracing/Assets/Shaders/Source/Ray/Test.hlsl:8,12-13 and
// there is no upload, hierarchy or material code to stand behind — SURVEY.md §0).
// Only argument checking, ownership and stream plumbing live here; the work is in
// scene_build.cu (GPU LBVH) and kernels_scene.cu (wavefront path tracer).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>

#include "ptb_internal.h"
#include "scene.cuh"

// ---- ptb_trace_rays_static: staging that is never allocated --------------------------------------------------------
// 32 B per ray: origins (12 B) | directions (12 B) | primitive ids (4 B) | distances (4 B), each part sized for the
// maximum batch so that every part starts on a 256 B boundary.  The array's name is a host-side stand-in in host code,
// not a device address: trace_static_scratch() asks the runtime for the address, once per device with that device
// current, and only that pointer is handed to the kernel.
#define PTB_TRACE_STATIC_BYTES ((size_t)PTB_TRACE_RAYS_STATIC_MAX * 32u)
static __device__ __align__(256) unsigned char g_trace_static_scratch[PTB_TRACE_STATIC_BYTES];

namespace {

using ptb::SceneState;

#define SC_CUDA(ctx, call)                                                                     \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess) return ptb::fail_cuda(ctx, e__, #call);                        \
    } while (0)

SceneState *state(ptb_context *ctx) { return static_cast<SceneState *>(ctx->scene); }

void free_queue(ptb::PathQueue &q)
{
    cudaFree(q.s0);
    cudaFree(q.s1);
    cudaFree(q.s2);
    q = ptb::PathQueue{};
}

void free_state(SceneState *st)
{
    if (!st) return;
    ptb::free_bvh(&st->sph_bvh);
    ptb::free_bvh(&st->tri_bvh);
    cudaFree(st->sph);
    cudaFree(st->sph_meta);
    cudaFree(st->tri);
    cudaFree(st->materials);
    free_queue(st->q[0]);
    free_queue(st->q[1]);
    cudaFree(st->counts);
    cudaFree(st->frame_accum);
    cudaFree(st->slices);
    cudaFree(st->stats);
    cudaFree(st->error_flag);
    delete st;
}

constexpr uint32_t MAX_WAVE_SAMPLES = 64u;                  // 6 bits of the path id word
// paths per wave the queues are sized for (2 x 40 B of path state + a 16 B radiance slot each = 6 GB at 64 Mi; measured
// 8 / 16 / 32 / 64 Mi-path waves: c3 3,326 / 3,425 / 3,919 / 3,966, c4 full frame 33 / 66 Mi: 1,112 / 1,130 Msamples/s)
constexpr uint32_t WAVE_PATHS_TARGET = 64u << 20;

ptb::SceneK make_k(const ptb_context *ctx, const SceneState *st)
{
    ptb::SceneK k{};
    k.width = ctx->cfg.width;
    k.height = ctx->cfg.height;
    k.row_offset = ctx->k.row_offset;
    k.row_stride = ctx->k.row_stride;
    k.row_block = ctx->k.row_block;
    k.div_x = ctx->cfg.uv_div_x;
    k.div_y = ctx->cfg.uv_div_y;
    for (int i = 0; i < 3; ++i) {
        k.cam_o[i] = st->desc.cam_origin[i];
        k.cam_u[i] = st->desc.cam_u[i];
        k.cam_v[i] = st->desc.cam_v[i];
        k.cam_w[i] = st->desc.cam_w[i];
        k.sky_h[i] = st->desc.sky_horizon[i];
        k.sky_z[i] = st->desc.sky_zenith[i];
    }
    k.t_min = st->desc.t_min;
    k.shade_mode = st->desc.shade_mode;
    k.max_depth = ctx->cfg.max_depth;
    k.seed = ctx->cfg.seed;
    k.n_spheres = st->desc.n_spheres;
    k.n_tris = st->desc.n_triangles;
    k.n_materials = st->desc.n_materials;
    k.sph = st->sph;
    k.sph_meta = st->sph_meta;
    k.sph_nodes = st->sph_bvh.nodes;
    k.sph_root = st->sph_bvh.root_ref;
    k.tri = st->tri;
    k.tri_nodes = st->tri_bvh.nodes;
    k.tri_root = st->tri_bvh.root_ref;
    k.materials = st->materials;
    k.n_sph_nodes = st->sph_bvh.n_nodes;
    k.n_tri_nodes = st->tri_bvh.n_nodes;
    k.sun_on = (st->sun_on && st->desc.shade_mode != PTB_SHADE_NORMALS) ? 1u : 0u;   // the normals mode has no light
    for (int i = 0; i < 3; ++i) {
        k.sun_dir[i] = st->sun_dir[i];
        k.sun_rad[i] = st->sun_rad[i];
    }
    return k;
}

uint32_t wave_paths(const ptb_context *ctx)
{
    return ptb::shard_row_count(ctx->cfg.height, ctx->k.row_offset, ctx->k.row_stride, ctx->k.row_block) * ctx->cfg.width;
}

// camera rays are indexed in 32 bits: 32 lanes per 8x4 pixel tile covering the row shard (times the samples of a wave)
int camera_ray_slots(ptb_context *ctx, uint64_t *slots)
{
    const uint64_t rows = ptb::shard_row_count(ctx->cfg.height, ctx->k.row_offset, ctx->k.row_stride, ctx->k.row_block);
    *slots = std::max<uint64_t>(1u, (((uint64_t)ctx->cfg.width + 7u) / 8u) * ((rows + 3u) / 4u) * 32u);
    if (*slots > 0xFFFFFFFFull)
        return ptb::fail_msg(ctx, PTB_ERR_UNSUPPORTED, "scene path: the image's 8x4 tile grid exceeds 2^32 camera-ray slots");
    return PTB_OK;
}

int check_error_flag(ptb_context *ctx, SceneState *st)
{
    uint32_t flag = 0;
    SC_CUDA(ctx, cudaMemcpyAsync(&flag, st->error_flag, sizeof flag, cudaMemcpyDeviceToHost, ctx->stream));
    SC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (flag & 1u)
        return ptb::fail_msg(ctx, PTB_ERR_UNSUPPORTED, "scene traversal stack overflow (hierarchy deeper than 64): results invalid");
    if (flag & ~1u) {                                   // only the -DPTB_BOUNDS_CHECK build sets these (kernels_scene.cu PTB_CHK)
        char msg[128];
        snprintf(msg, sizeof msg, "scene kernels: bounds check failed, flags 0x%x (see PTB_CHK in kernels_scene.cu)", flag);
        return ptb::fail_msg(ctx, PTB_ERR_CUDA, msg);
    }
    return PTB_OK;
}

// Device address of g_trace_static_scratch on the context's device (which is current): looked up once per device.
constexpr int MAX_STATIC_DEVICES = 64;
unsigned char *g_trace_static_addr[MAX_STATIC_DEVICES] = {};
std::mutex g_trace_static_mutex;

int trace_static_scratch(ptb_context *ctx, unsigned char **out)
{
    const int dev = ctx->cfg.device;
    if (dev < 0 || dev >= MAX_STATIC_DEVICES)
        return ptb::fail_msg(ctx, PTB_ERR_UNSUPPORTED, "ptb_trace_rays_static: device index beyond the address table");
    std::lock_guard<std::mutex> lock(g_trace_static_mutex);
    if (!g_trace_static_addr[dev]) {
        void *p = nullptr;
        SC_CUDA(ctx, cudaGetSymbolAddress(&p, g_trace_static_scratch));
        if (!p) return ptb::fail_msg(ctx, PTB_ERR_CUDA, "ptb_trace_rays_static: cudaGetSymbolAddress returned NULL");
        g_trace_static_addr[dev] = static_cast<unsigned char *>(p);
    }
    *out = g_trace_static_addr[dev];
    return PTB_OK;
}

}  // namespace

namespace ptb {

void scene_destroy(ptb_context *ctx)
{
    free_state(static_cast<SceneState *>(ctx->scene));
    ctx->scene = nullptr;
}

bool scene_resident(const ptb_context *ctx) { return ctx->scene != nullptr; }

int scene_compute_frame(ptb_context *ctx)
{
    SceneState *st = state(ctx);
    const size_t n = (size_t)ctx->cfg.width * ctx->cfg.height;
    uint64_t slots = 0;
    if (int rc = camera_ray_slots(ctx, &slots)) return rc;
    if (!st->frame_accum) SC_CUDA(ctx, cudaMalloc(&st->frame_accum, n * sizeof(float4)));
    SC_CUDA(ctx, cudaMemsetAsync(st->frame_accum, 0, n * sizeof(float4), ctx->stream));
    if (ctx->buf.hit) SC_CUDA(ctx, cudaMemsetAsync(ctx->buf.hit, 0xFF, n * sizeof(int32_t), ctx->stream));
    const SceneK k = make_k(ctx, st);
    st->last_n0 = wave_paths(ctx);
    SC_CUDA(ctx, launch_scene_wave(k, st, st->frame_accum, ctx->buf.hit, 0u, 1u, false, ctx->stream, &ctx->launches));
    SC_CUDA(ctx, launch_frame_finish(st->frame_accum, ctx->buf.f32, ctx->buf.rgba8, (uint32_t)n, ctx->stream));
    ctx->launches += 1;
    return PTB_OK;
}

int scene_render_samples(ptb_context *ctx, uint32_t first_sample, uint32_t n_samples, int jitter)
{
    SceneState *st = state(ctx);
    const SceneK k = make_k(ctx, st);
    // One wave traces as many samples of every pixel of the shard as the path queues hold: small frames and 1/N row shards
    // would otherwise run waves too small to fill the GPU at the deep bounces (profiles/r01_notes.md).  Same results bit
    // for bit: the samples land in per-sample slices that are added to the accumulator in sample order.
    const uint32_t shard = wave_paths(ctx);
    if (shard == 0 || n_samples == 0) return PTB_OK;
    uint64_t slots = 0;
    if (int rc = camera_ray_slots(ctx, &slots)) return rc;
    uint32_t per_wave = std::min<uint32_t>(MAX_WAVE_SAMPLES, std::max<uint32_t>(1u, st->q_cap / shard));
    if (st->wave_samples) per_wave = std::min(per_wave, st->wave_samples);
    if (shard > (1u << 26)) per_wave = 1u;              // a batched wave keeps 26 bits of the path id word for the pixel
    per_wave = std::min(per_wave, n_samples);
    per_wave = (uint32_t)std::max<uint64_t>(1u, std::min<uint64_t>(per_wave, 0xFFFFFFFFull / slots));   // 32-bit camera-ray index
    if (per_wave > 1u && !st->slices)                   // one radiance slot per path of a full wave
        SC_CUDA(ctx, cudaMalloc(&st->slices, (size_t)st->q_cap * sizeof(float4)));
    for (uint32_t s = 0; s < n_samples; s += per_wave) {
        const uint32_t batch = std::min(per_wave, n_samples - s);
        st->last_n0 = shard * batch;
        SC_CUDA(ctx, launch_scene_wave(k, st, ctx->buf.accum, nullptr, first_sample + s, batch, jitter != 0, ctx->stream, &ctx->launches));
    }
    return PTB_OK;
}

int scene_wait_check(ptb_context *ctx) { return check_error_flag(ctx, state(ctx)); }

}  // namespace ptb

extern "C" {

int ptb_clear_scene(ptb_context *ctx)
{
    if (!ctx) return ptb::fail_msg(nullptr, PTB_ERR_INVALID, "ptb_clear_scene: ctx is NULL");
    SC_CUDA(ctx, cudaSetDevice(ctx->cfg.device));
    SC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ptb::scene_destroy(ctx);
    return PTB_OK;
}

int ptb_upload_scene(ptb_context *ctx, const ptb_scene_desc *d)
{
    if (!ctx) return ptb::fail_msg(nullptr, PTB_ERR_INVALID, "ptb_upload_scene: ctx is NULL");
    if (!d) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_upload_scene: desc is NULL");
    if (d->struct_size != sizeof(ptb_scene_desc))
        return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_upload_scene: struct_size mismatch (ABI)");
    if (d->n_spheres >= ptb::MAX_PRIMS_PER_CLASS || d->n_triangles >= ptb::MAX_PRIMS_PER_CLASS)
        return ptb::fail_msg(ctx, PTB_ERR_UNSUPPORTED, "ptb_upload_scene: more than 2^27 primitives of one class");
    if ((d->n_spheres && (!d->spheres || !d->sphere_material)) || (d->n_triangles && (!d->tri_vertices || !d->tri_material)))
        return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_upload_scene: primitive array is NULL");
    if (d->n_materials == 0 || !d->materials) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_upload_scene: at least one material is required");
    if (d->shade_mode > PTB_SHADE_NORMALS) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_upload_scene: unknown shade_mode");
    if (d->max_leaf > ptb::MAX_LEAF) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_upload_scene: max_leaf > 8");
    if (d->flags & ~PTB_SCENE_COUNT_STATS) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_upload_scene: unknown flags");
    if (!(d->t_min >= 0.0f)) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_upload_scene: t_min must be >= 0");
    for (uint32_t i = 0; i < d->n_materials; ++i)
        if (d->materials[i].type > PTB_MAT_DIELECTRIC) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_upload_scene: unknown material type");
    SC_CUDA(ctx, cudaSetDevice(ctx->cfg.device));
    SC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ptb::scene_destroy(ctx);

    // largest coordinate magnitude (host pass): the boxes are padded by 2^-18 of it so that
    // culling stays conservative against the rounding of the primitive tests
    // (std::fmax drops NaN operands, so finiteness is tested on every value, not on the running maximum)
    float ext = 0.0f;
    bool finite = std::isfinite(d->t_min);
    for (int i = 0; i < 3; ++i) {
        finite = finite && std::isfinite(d->cam_origin[i]) && std::isfinite(d->cam_u[i]) && std::isfinite(d->cam_v[i]) &&
                 std::isfinite(d->cam_w[i]);
        ext = std::fmax(ext, std::fabs(d->cam_origin[i]));
    }
    for (uint32_t i = 0; i < d->n_spheres; ++i) {
        const float *c = d->spheres + 4 * (size_t)i;
        const float r = std::fabs(c[3]);
        finite = finite && std::isfinite(c[0]) && std::isfinite(c[1]) && std::isfinite(c[2]) && std::isfinite(c[3]);
        for (int kx = 0; kx < 3; ++kx) ext = std::fmax(ext, std::fabs(c[kx]) + r);
    }
    for (size_t i = 0; i < (size_t)d->n_triangles * 9; ++i) {
        finite = finite && std::isfinite(d->tri_vertices[i]);
        ext = std::fmax(ext, std::fabs(d->tri_vertices[i]));
    }
    if (!finite || !std::isfinite(ext)) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_upload_scene: non-finite coordinate");
    const float pad = ext * 3.814697265625e-6f;

    SceneState *st = new (std::nothrow) SceneState();
    if (!st) return ptb::fail_msg(ctx, PTB_ERR_NOMEM, "ptb_upload_scene: host allocation failed");
    ctx->scene = st;
    st->desc = *d;
    st->desc.spheres = nullptr; st->desc.sphere_material = nullptr; st->desc.tri_vertices = nullptr;
    st->desc.tri_material = nullptr; st->desc.materials = nullptr;
    st->count_stats = (d->flags & PTB_SCENE_COUNT_STATS) != 0;
    const uint32_t max_leaf = d->max_leaf ? d->max_leaf : 4u;
    const uint32_t ns = d->n_spheres, nt = d->n_triangles;

    float4 *raw_sph = nullptr, *lo = nullptr, *hi = nullptr;
    uint32_t *raw_smat = nullptr, *raw_tmat = nullptr;
    float *raw_tri = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    int rc = PTB_OK;
    auto cleanup = [&]() {
        cudaFree(raw_sph); cudaFree(lo); cudaFree(hi); cudaFree(raw_smat); cudaFree(raw_tmat); cudaFree(raw_tri);
        if (e0) cudaEventDestroy(e0);
        if (e1) cudaEventDestroy(e1);
    };
#define UP(call)                                                   \
    do {                                                           \
        cudaError_t e__ = (call);                                  \
        if (e__ != cudaSuccess) {                                  \
            rc = ptb::fail_cuda(ctx, e__, #call);                  \
            cleanup();                                             \
            ptb::scene_destroy(ctx);                               \
            return rc;                                             \
        }                                                          \
    } while (0)

    const uint32_t nmax = ns > nt ? ns : nt;
    UP(cudaEventCreate(&e0));
    UP(cudaEventCreate(&e1));
    UP(cudaMalloc(&lo, sizeof(float4) * (nmax ? nmax : 1)));
    UP(cudaMalloc(&hi, sizeof(float4) * (nmax ? nmax : 1)));
    if (ns) {
        UP(cudaMalloc(&raw_sph, sizeof(float4) * ns));
        UP(cudaMalloc(&raw_smat, sizeof(uint32_t) * ns));
        UP(cudaMalloc(&st->sph, sizeof(float4) * ns));
        UP(cudaMalloc(&st->sph_meta, sizeof(uint2) * ns));
        UP(cudaMemcpyAsync(raw_sph, d->spheres, sizeof(float4) * ns, cudaMemcpyHostToDevice, ctx->stream));
        UP(cudaMemcpyAsync(raw_smat, d->sphere_material, sizeof(uint32_t) * ns, cudaMemcpyHostToDevice, ctx->stream));
    }
    if (nt) {
        UP(cudaMalloc(&raw_tri, sizeof(float) * 9 * (size_t)nt));
        UP(cudaMalloc(&raw_tmat, sizeof(uint32_t) * nt));
        UP(cudaMalloc(&st->tri, sizeof(float4) * 3 * (size_t)nt));
        UP(cudaMemcpyAsync(raw_tri, d->tri_vertices, sizeof(float) * 9 * (size_t)nt, cudaMemcpyHostToDevice, ctx->stream));
        UP(cudaMemcpyAsync(raw_tmat, d->tri_material, sizeof(uint32_t) * nt, cudaMemcpyHostToDevice, ctx->stream));
    }
    UP(cudaMalloc(&st->materials, sizeof(ptb_material) * d->n_materials));
    UP(cudaMemcpyAsync(st->materials, d->materials, sizeof(ptb_material) * d->n_materials, cudaMemcpyHostToDevice, ctx->stream));

    UP(cudaEventRecord(e0, ctx->stream));
    st->sph_bvh.root_ref = ptb::EMPTY_REF;
    st->tri_bvh.root_ref = ptb::EMPTY_REF;
    if (ns) {
        UP(ptb::launch_sphere_bounds(raw_sph, ns, lo, hi, ctx->stream));
        UP(ptb::build_lbvh(lo, hi, ns, max_leaf, pad, &st->sph_bvh, ctx->stream));
        UP(ptb::launch_pack_spheres(raw_sph, raw_smat, st->sph_bvh.order, ns, st->sph, st->sph_meta, ctx->stream));
    }
    if (nt) {
        UP(ptb::launch_tri_bounds(raw_tri, nt, lo, hi, ctx->stream));
        UP(ptb::build_lbvh(lo, hi, nt, max_leaf, pad, &st->tri_bvh, ctx->stream));
        UP(ptb::launch_pack_tris(raw_tri, raw_tmat, st->tri_bvh.order, nt, ns, st->tri, ctx->stream));
    }
    UP(cudaEventRecord(e1, ctx->stream));
    UP(cudaEventSynchronize(e1));
    UP(cudaEventElapsedTime(&st->info.build_ms, e0, e1));
    ctx->launches += 12;     // bounds, Morton, sort passes, Karras, refit, survive, scan, emit, pack (per class; lower bound)

    // wavefront state
    {
        // wave capacity: at least one sample of every pixel; small frames get room for several samples per wave
        // (PTB200_WAVE_PATHS overrides the target for A/B runs)
        static const uint64_t target = [] {
            const char *e = std::getenv("PTB200_WAVE_PATHS");
            return e ? (uint64_t)std::strtoull(e, nullptr, 10) : (uint64_t)WAVE_PATHS_TARGET;
        }();
        const uint64_t n_pix = (uint64_t)ctx->cfg.width * ctx->cfg.height;
        st->q_cap = (uint32_t)std::max<uint64_t>(n_pix, std::min<uint64_t>(target, n_pix * MAX_WAVE_SAMPLES));
    }
    for (int i = 0; i < 2; ++i) {
        UP(cudaMalloc(&st->q[i].s0, sizeof(float4) * st->q_cap));
        UP(cudaMalloc(&st->q[i].s1, sizeof(float4) * st->q_cap));
        UP(cudaMalloc(&st->q[i].s2, sizeof(float2) * st->q_cap));
    }
    // one radiance slot per path of a full wave (batched waves).  Allocated here rather than at the first render: cudaMalloc
    // synchronises the device, which inside a pipelined multi-GPU frame loop would stall behind a rendezvous kernel
    UP(cudaMalloc(&st->slices, (size_t)st->q_cap * sizeof(float4)));
    st->counts_len = 2u * (ctx->cfg.max_depth + 2u);
    UP(cudaMalloc(&st->counts, sizeof(uint32_t) * st->counts_len));
    UP(cudaMalloc(&st->stats, sizeof(ptb::SceneStatsDev)));
    UP(cudaMemsetAsync(st->stats, 0, sizeof(ptb::SceneStatsDev), ctx->stream));
    UP(cudaMalloc(&st->error_flag, sizeof(uint32_t)));
    UP(cudaMemsetAsync(st->error_flag, 0, sizeof(uint32_t), ctx->stream));
    UP(cudaStreamSynchronize(ctx->stream));
#undef UP
    cleanup();

    st->info.n_spheres = ns;
    st->info.n_triangles = nt;
    st->info.sphere_nodes = st->sph_bvh.n_nodes;
    st->info.triangle_nodes = st->tri_bvh.n_nodes;
    st->info.sphere_leaves = st->sph_bvh.n_leaves;
    st->info.triangle_leaves = st->tri_bvh.n_leaves;
    st->info.max_leaf = max_leaf;
    st->info.device_bytes = (uint64_t)st->sph_bvh.n_nodes * 64 + (uint64_t)st->tri_bvh.n_nodes * 64 + (uint64_t)ns * 24 +
                            (uint64_t)nt * 48 + (uint64_t)d->n_materials * sizeof(ptb_material);
    return PTB_OK;
}

int ptb_scene_info_get(ptb_context *ctx, ptb_scene_info *out)
{
    if (!ctx) return ptb::fail_msg(nullptr, PTB_ERR_INVALID, "ptb_scene_info_get: ctx is NULL");
    if (!out) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_scene_info_get: out is NULL");
    if (!ctx->scene) return ptb::fail_msg(ctx, PTB_ERR_STATE, "ptb_scene_info_get: no scene resident");
    *out = state(ctx)->info;
    return PTB_OK;
}

int ptb_scene_stats_get(ptb_context *ctx, ptb_scene_stats *out)
{
    if (!ctx) return ptb::fail_msg(nullptr, PTB_ERR_INVALID, "ptb_scene_stats_get: ctx is NULL");
    if (!out) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_scene_stats_get: out is NULL");
    if (!ctx->scene) return ptb::fail_msg(ctx, PTB_ERR_STATE, "ptb_scene_stats_get: no scene resident");
    SceneState *st = state(ctx);
    if (!st->count_stats) return ptb::fail_msg(ctx, PTB_ERR_STATE, "ptb_scene_stats_get: scene uploaded without PTB_SCENE_COUNT_STATS");
    SC_CUDA(ctx, cudaSetDevice(ctx->cfg.device));
    ptb::SceneStatsDev h{};
    SC_CUDA(ctx, cudaMemcpyAsync(&h, st->stats, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
    SC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    out->rays = h.rays;
    out->nodes_visited = h.nodes_visited;
    out->prims_tested = h.prims_tested;
    out->scatters = h.scatters;
    out->paths = h.paths;
    out->max_stack = h.max_stack;
    return check_error_flag(ctx, st);
}

int ptb_scene_stats_reset(ptb_context *ctx)
{
    if (!ctx) return ptb::fail_msg(nullptr, PTB_ERR_INVALID, "ptb_scene_stats_reset: ctx is NULL");
    if (!ctx->scene) return ptb::fail_msg(ctx, PTB_ERR_STATE, "ptb_scene_stats_reset: no scene resident");
    SC_CUDA(ctx, cudaSetDevice(ctx->cfg.device));
    SC_CUDA(ctx, cudaMemsetAsync(state(ctx)->stats, 0, sizeof(ptb::SceneStatsDev), ctx->stream));
    return PTB_OK;
}

int ptb_scene_set_wave_samples(ptb_context *ctx, uint32_t samples_per_wave)
{
    if (!ctx) return ptb::fail_msg(nullptr, PTB_ERR_INVALID, "ptb_scene_set_wave_samples: ctx is NULL");
    if (!ctx->scene) return ptb::fail_msg(ctx, PTB_ERR_STATE, "ptb_scene_set_wave_samples: no scene resident");
    state(ctx)->wave_samples = samples_per_wave;
    return PTB_OK;
}

int ptb_scene_set_sun(ptb_context *ctx, const float dir_to_light[3], const float radiance[3])
{
    if (!ctx) return ptb::fail_msg(nullptr, PTB_ERR_INVALID, "ptb_scene_set_sun: ctx is NULL");
    if (!dir_to_light || !radiance) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_scene_set_sun: NULL argument");
    if (!ctx->scene) return ptb::fail_msg(ctx, PTB_ERR_STATE, "ptb_scene_set_sun: no scene resident");
    bool finite = true, dir_nonzero = false, on = false;
    for (int i = 0; i < 3; ++i) {
        finite = finite && std::isfinite(dir_to_light[i]) && std::isfinite(radiance[i]);
        dir_nonzero = dir_nonzero || dir_to_light[i] != 0.0f;
        on = on || radiance[i] != 0.0f;
    }
    if (!finite) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_scene_set_sun: direction and radiance must be finite");
    if (on && !dir_nonzero) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_scene_set_sun: the direction is the zero vector");
    SceneState *st = state(ctx);
    st->sun_on = on;
    for (int i = 0; i < 3; ++i) {
        st->sun_dir[i] = dir_to_light[i];
        st->sun_rad[i] = radiance[i];
    }
    return PTB_OK;
}

int ptb_scene_queue_sizes(ptb_context *ctx, uint32_t *out, uint32_t cap, uint32_t *n)
{
    if (!ctx) return ptb::fail_msg(nullptr, PTB_ERR_INVALID, "ptb_scene_queue_sizes: ctx is NULL");
    if (!out || !n) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_scene_queue_sizes: NULL argument");
    if (!ctx->scene) return ptb::fail_msg(ctx, PTB_ERR_STATE, "ptb_scene_queue_sizes: no scene resident");
    SC_CUDA(ctx, cudaSetDevice(ctx->cfg.device));
    SceneState *st = state(ctx);
    uint32_t cnt = ctx->cfg.max_depth + 1u;
    if (cnt > cap) cnt = cap;
    if (cnt) SC_CUDA(ctx, cudaMemcpyAsync(out, st->counts, sizeof(uint32_t) * cnt, cudaMemcpyDeviceToHost, ctx->stream));
    SC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (cnt) out[0] = st->last_n0;
    *n = cnt;
    return PTB_OK;
}

int ptb_scene_read_bvh(ptb_context *ctx, int which, void *nodes_out, uint32_t cap_nodes, uint32_t *n_nodes, uint32_t *order_out,
                       uint32_t cap_prims, uint32_t *n_prims)
{
    if (!ctx) return ptb::fail_msg(nullptr, PTB_ERR_INVALID, "ptb_scene_read_bvh: ctx is NULL");
    if (!ctx->scene) return ptb::fail_msg(ctx, PTB_ERR_STATE, "ptb_scene_read_bvh: no scene resident");
    if (which < 0 || which > 1) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_scene_read_bvh: which must be 0 or 1");
    SC_CUDA(ctx, cudaSetDevice(ctx->cfg.device));
    const ptb::DeviceBvh &b = which == 0 ? state(ctx)->sph_bvh : state(ctx)->tri_bvh;
    if (n_nodes) *n_nodes = b.n_nodes;
    if (n_prims) *n_prims = b.n_prims;
    if (nodes_out) {
        if (cap_nodes < b.n_nodes) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_scene_read_bvh: node buffer too small");
        if (b.n_nodes) SC_CUDA(ctx, cudaMemcpyAsync(nodes_out, b.nodes, (size_t)b.n_nodes * 64, cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (order_out) {
        if (cap_prims < b.n_prims) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_scene_read_bvh: order buffer too small");
        if (b.n_prims) SC_CUDA(ctx, cudaMemcpyAsync(order_out, b.order, (size_t)b.n_prims * 4, cudaMemcpyDeviceToHost, ctx->stream));
    }
    SC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PTB_OK;
}

int ptb_trace_rays(ptb_context *ctx, uint64_t n, const float *origins, const float *dirs, float t_min, float t_max, int mode,
                   int32_t *out_prim, float *out_t)
{
    if (!ctx) return ptb::fail_msg(nullptr, PTB_ERR_INVALID, "ptb_trace_rays: ctx is NULL");
    if (!ctx->scene) return ptb::fail_msg(ctx, PTB_ERR_STATE, "ptb_trace_rays: no scene resident");
    if (mode < 0 || mode > 1) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_trace_rays: mode must be 0 (closest) or 1 (any)");
    if (n == 0) return PTB_OK;
    if (!origins || !dirs || !out_prim) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_trace_rays: NULL array");
    if (n > (1ull << 31)) return ptb::fail_msg(ctx, PTB_ERR_UNSUPPORTED, "ptb_trace_rays: more than 2^31 rays");
    if (!(t_min >= 0.0f) || !(t_max >= t_min))
        return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_trace_rays: need 0 <= t_min <= t_max (the conservative culling interval is "
                             "widened relative to non-negative distances)");
    SC_CUDA(ctx, cudaSetDevice(ctx->cfg.device));
    SceneState *st = state(ctx);
    float *d_o = nullptr, *d_d = nullptr, *d_t = nullptr;
    int32_t *d_p = nullptr;
    int rc = PTB_OK;
    cudaError_t e = cudaSuccess;
    do {
        if ((e = cudaMalloc(&d_o, n * 12)) != cudaSuccess) break;
        if ((e = cudaMalloc(&d_d, n * 12)) != cudaSuccess) break;
        if ((e = cudaMalloc(&d_p, n * 4)) != cudaSuccess) break;
        if ((e = cudaMalloc(&d_t, n * 4)) != cudaSuccess) break;
        if ((e = cudaMemcpyAsync(d_o, origins, n * 12, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) break;
        if ((e = cudaMemcpyAsync(d_d, dirs, n * 12, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) break;
        const ptb::SceneK k = make_k(ctx, st);
        if ((e = ptb::launch_trace_rays(k, st, n, d_o, d_d, t_min, t_max, mode, d_p, d_t, ctx->stream)) != cudaSuccess) break;
        ctx->launches += 1;
        if ((e = cudaMemcpyAsync(out_prim, d_p, n * 4, cudaMemcpyDeviceToHost, ctx->stream)) != cudaSuccess) break;
        if (out_t && mode == 0)
            if ((e = cudaMemcpyAsync(out_t, d_t, n * 4, cudaMemcpyDeviceToHost, ctx->stream)) != cudaSuccess) break;
        e = cudaStreamSynchronize(ctx->stream);
    } while (0);
    cudaFree(d_o); cudaFree(d_d); cudaFree(d_p); cudaFree(d_t);
    if (e != cudaSuccess) return ptb::fail_cuda(ctx, e, "ptb_trace_rays");
    rc = check_error_flag(ctx, st);
    return rc;
}

int ptb_trace_rays_static(ptb_context *ctx, uint64_t n, const float *origins, const float *dirs, float t_min, float t_max, int mode,
                          int32_t *out_prim, float *out_t)
{
    if (!ctx) return ptb::fail_msg(nullptr, PTB_ERR_INVALID, "ptb_trace_rays_static: ctx is NULL");
    if (!ctx->scene) return ptb::fail_msg(ctx, PTB_ERR_STATE, "ptb_trace_rays_static: no scene resident");
    if (mode < 0 || mode > 1) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_trace_rays_static: mode must be 0 (closest) or 1 (any)");
    if (n == 0) return PTB_OK;
    if (!origins || !dirs || !out_prim) return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_trace_rays_static: NULL array");
    if (n > PTB_TRACE_RAYS_STATIC_MAX) {
        char msg[160];
        snprintf(msg, sizeof msg, "ptb_trace_rays_static: %llu rays exceed PTB_TRACE_RAYS_STATIC_MAX = %u (the static staging "
                 "array); split the batch or use ptb_trace_rays", (unsigned long long)n, PTB_TRACE_RAYS_STATIC_MAX);
        return ptb::fail_msg(ctx, PTB_ERR_UNSUPPORTED, msg);
    }
    if (!(t_min >= 0.0f) || !(t_max >= t_min))
        return ptb::fail_msg(ctx, PTB_ERR_INVALID, "ptb_trace_rays_static: need 0 <= t_min <= t_max (the conservative culling "
                             "interval is widened relative to non-negative distances)");
    SC_CUDA(ctx, cudaSetDevice(ctx->cfg.device));
    unsigned char *base = nullptr;
    const int rc = trace_static_scratch(ctx, &base);
    if (rc != PTB_OK) return rc;
    SceneState *st = state(ctx);
    const size_t cap = PTB_TRACE_RAYS_STATIC_MAX;
    float *d_o = reinterpret_cast<float *>(base);
    float *d_d = reinterpret_cast<float *>(base + cap * 12);
    int32_t *d_p = reinterpret_cast<int32_t *>(base + cap * 24);
    float *d_t = reinterpret_cast<float *>(base + cap * 28);
    // the array keeps what the last call left: the results start from zero, the ray parts are overwritten up to n
    SC_CUDA(ctx, cudaMemsetAsync(d_p, 0, n * 4, ctx->stream));
    SC_CUDA(ctx, cudaMemsetAsync(d_t, 0, n * 4, ctx->stream));
    SC_CUDA(ctx, cudaMemcpyAsync(d_o, origins, n * 12, cudaMemcpyHostToDevice, ctx->stream));
    SC_CUDA(ctx, cudaMemcpyAsync(d_d, dirs, n * 12, cudaMemcpyHostToDevice, ctx->stream));
    const ptb::SceneK k = make_k(ctx, st);
    SC_CUDA(ctx, ptb::launch_trace_rays(k, st, n, d_o, d_d, t_min, t_max, mode, d_p, d_t, ctx->stream));
    ctx->launches += 1;
    SC_CUDA(ctx, cudaMemcpyAsync(out_prim, d_p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_t && mode == 0) SC_CUDA(ctx, cudaMemcpyAsync(out_t, d_t, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    SC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return check_error_flag(ctx, st);
}

}  // extern "C"
