// api.cu — the extern "C" boundary of libptb200.so (include/ptb200.h).
// Each entry point names the reference member it replaces in the header; this file only does
// argument checking, resource ownership and stream/event plumbing around the kernel launchers.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <utility>
#include <new>

#include "ptb_internal.h"

namespace {

thread_local std::string g_err;   // errors with no context to hang them on

int fail(ptb_context *ctx, int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf;
    else g_err = buf;
    return code;
}

int cuda_fail(ptb_context *ctx, cudaError_t e, const char *what)
{
    return fail(ctx, e == cudaErrorMemoryAllocation ? PTB_ERR_NOMEM : PTB_ERR_CUDA, "%s: %s (%s)", what,
                cudaGetErrorString(e), cudaGetErrorName(e));
}

#define PTB_CUDA(ctx, call)                                        \
    do {                                                           \
        cudaError_t e__ = (call);                                  \
        if (e__ != cudaSuccess) return cuda_fail(ctx, e__, #call); \
    } while (0)

// Every entry point makes the context's device current (contexts on different devices may be
// interleaved on one host thread).
#define PTB_ENTER(ctx)                                                    \
    do {                                                                  \
        if (!(ctx)) return fail(nullptr, PTB_ERR_INVALID, "%s: ctx is NULL", __func__); \
        PTB_CUDA(ctx, cudaSetDevice((ctx)->cfg.device));                  \
    } while (0)

size_t n_pixels(const ptb_context *c) { return (size_t)c->cfg.width * c->cfg.height; }

void free_buffers(ptb_context *c)
{
    cudaFree(c->buf.rgba8);
    cudaFree(c->buf.rgba8_alt);
    cudaFree(c->buf.f32);
    cudaFree(c->buf.hit);
    cudaFree(c->buf.accum_owned);
    cudaFree(c->buf.present);
    cudaFree(c->buf.row_span);
    c->buf = ptb::Buffers{};
}

int read_rows(ptb_context *ctx, void *host_dst, size_t pitch, const void *dev, size_t row_bytes, size_t rows)
{
    if (!host_dst) return fail(ctx, PTB_ERR_INVALID, "host destination is NULL");
    if (pitch == 0) pitch = row_bytes;
    if (pitch < row_bytes) return fail(ctx, PTB_ERR_INVALID, "pitch %zu < row size %zu", pitch, row_bytes);
    PTB_CUDA(ctx, cudaMemcpy2DAsync(host_dst, pitch, dev, row_bytes, row_bytes, rows, cudaMemcpyDeviceToHost,
                                    ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PTB_OK;
}

// Launch-uniform validity of the shortened exact sequences (kernels_csmain.cu MODE_FAST*):
//  - both screen divisors pass the exhaustive one-correction division check for v in [2^-24, 2^17);
//  - pixel coordinates (plus jitter) stay below 2^17;
//  - |uv| is bounded so that dot(v,v) of Test.hlsl:10 stays inside [1, 2^23) (sqrt / rcp domains);
//  - every derived term is finite.
void derive_fast_flags(ptb_context *c)
{
    ptb::FrameK &k = c->k;
    k.fast = 0;
    k.on_axis = 0;
    k.cm_on = 0;
    k.cm_sx = k.cm_sy = k.cm_K = 0.0f;
    k.div2 = 0;
    k.one = 1u;
    k.dl_x = k.dl_y = 0.0f;
    k.spirv = (c->cfg.flags & PTB_CFG_SPIRV_ARITH) ? 1u : 0u;
    if (c->cfg.flags & (PTB_CFG_FORCE_IEEE | PTB_CFG_SPIRV_ARITH)) return;
    const float vals[] = {k.cam[0], k.cam[1], k.cam[2], k.centre[0], k.centre[1], k.centre[2], k.radius,
                          k.oc[0], k.oc[1], k.oc[2], k.c, k.c4, k.inv_div_x, k.inv_div_y};
    for (float v : vals)
        if (!std::isfinite(v)) return;
    if (k.width > 65536u || k.height > 65536u) return;
    const double ux = 2.0 * (k.width + 1.0) / std::fabs((double)k.div_x) + 1.0;
    const double uy = 2.0 * (k.height + 1.0) / std::fabs((double)k.div_y) + 1.0;
    if (!(ux * ux + uy * uy + 1.0 < 8388608.0)) return;
    const int vx = ptb::uniform_divisor_verified(k.div_x, c->stream), vy = ptb::uniform_divisor_verified(k.div_y, c->stream);
    if (!(vx & 1) || !(vy & 1)) return;
    k.fast = 1;
    k.div2 = ((vx & 2) && (vy & 2) && !std::getenv("PTB200_NO_DIV2")) ? 1u : 0u;
    k.dl_x = ptb::reciprocal_low(k.div_x);
    k.dl_y = ptb::reciprocal_low(k.div_y);
    // per-pixel certain-miss classification: only bounds are needed, so plain host arithmetic is fine
    k.cm_sx = 2.0f / k.div_x;
    k.cm_sy = 2.0f / k.div_y;
    k.cm_K = std::fabs(k.c) + (k.oc[0] * k.oc[0] + k.oc[1] * k.oc[1] + k.oc[2] * k.oc[2]);
    k.cm_on = (std::isfinite(k.cm_K) && k.cm_K > 0.0f && k.cm_K < 1e30f && !std::getenv("PTB200_NO_CERTAIN_MISS")) ? 1u : 0u;
    const float az = std::fabs(k.oc[2]);
    k.on_axis = (k.oc[0] == 0.0f && k.oc[1] == 0.0f && az >= 8.67e-19f /*~2^-60*/ && az <= 1.15e18f /*~2^60*/) ? 1u : 0u;
}

}  // namespace

namespace ptb {
// shared with scene.cu (declared in ptb_internal.h)
int fail_msg(ptb_context *ctx, int code, const char *msg) { return fail(ctx, code, "%s", msg); }
int fail_cuda(ptb_context *ctx, cudaError_t e, const char *what) { return cuda_fail(ctx, e, what); }
}

extern "C" {

int ptb_abi_version(void) { return PTB_ABI_VERSION; }

int ptb_device_count(void)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaGetDeviceCount");
    return n;
}

int ptb_default_config(ptb_config *cfg, uint32_t width, uint32_t height)
{
    if (!cfg) return fail(nullptr, PTB_ERR_INVALID, "ptb_default_config: cfg is NULL");
    std::memset(cfg, 0, sizeof *cfg);
    cfg->struct_size = (uint32_t)sizeof *cfg;
    cfg->device = 0;
    cfg->width = width;
    cfg->height = height;
    cfg->uv_div_x = (float)height;   // Test.hlsl:6 uses 1080 on both axes of a 1920x1080 image
    cfg->uv_div_y = (float)height;
    cfg->cam[2] = 1.0f;              // Test.hlsl:8
    cfg->radius = 0.5f;              // Test.hlsl:13 (centre (0,0,0), :12)
    return PTB_OK;
}

int ptb_create(const ptb_config *cfg, ptb_context **out_ctx)
{
    if (out_ctx) *out_ctx = nullptr;
    if (!cfg || !out_ctx) return fail(nullptr, PTB_ERR_INVALID, "ptb_create: NULL argument");
    if (cfg->struct_size != sizeof(ptb_config))
        return fail(nullptr, PTB_ERR_INVALID, "ptb_create: struct_size %u != %zu (ABI mismatch)", cfg->struct_size,
                    sizeof(ptb_config));
    if (cfg->width == 0 || cfg->height == 0)
        return fail(nullptr, PTB_ERR_INVALID, "ptb_create: empty image %ux%u", cfg->width, cfg->height);
    if ((uint64_t)cfg->width * cfg->height > (1ull << 30))
        return fail(nullptr, PTB_ERR_UNSUPPORTED, "ptb_create: %ux%u exceeds 2^30 pixels", cfg->width, cfg->height);
    if (!(cfg->uv_div_x != 0.0f) || !(cfg->uv_div_y != 0.0f))
        return fail(nullptr, PTB_ERR_INVALID, "ptb_create: uv divisor must be non-zero");
    if (cfg->flags & ~(PTB_CFG_KEEP_F32 | PTB_CFG_KEEP_HIT_IDS | PTB_CFG_EXTERNAL_STREAM | PTB_CFG_FORCE_IEEE | PTB_CFG_SPIRV_ARITH))
        return fail(nullptr, PTB_ERR_INVALID, "ptb_create: unknown flags 0x%x", cfg->flags);
    if ((cfg->flags & PTB_CFG_SPIRV_ARITH) &&
        !(cfg->cam[0] == 0.0f && cfg->cam[1] == 0.0f && cfg->cam[2] == 1.0f && cfg->centre[0] == 0.0f && cfg->centre[1] == 0.0f &&
          cfg->centre[2] == 0.0f && cfg->radius == 0.5f))
        return fail(nullptr, PTB_ERR_UNSUPPORTED, "ptb_create: PTB_CFG_SPIRV_ARITH follows the compiled module, which has the "
                    "literal camera (0,0,1) and sphere (0,0,0; 0.5) of Test.hlsl:8,12-13 folded in");
    if (cfg->max_depth > 1024u)
        return fail(nullptr, PTB_ERR_INVALID, "ptb_create: max_depth %u exceeds the limit of 1024", cfg->max_depth);

    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "ptb_create: no usable CUDA driver/device (no CPU fallback)");
    if (cfg->device < 0 || cfg->device >= ndev)
        return fail(nullptr, PTB_ERR_INVALID, "ptb_create: device %d out of range [0,%d)", cfg->device, ndev);
    e = cudaSetDevice(cfg->device);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaSetDevice");

    ptb_context *c = new (std::nothrow) ptb_context();
    if (!c) return fail(nullptr, PTB_ERR_NOMEM, "ptb_create: host allocation failed");
    c->cfg = *cfg;
    c->k.width = cfg->width;
    c->k.height = cfg->height;
    c->k.row_offset = 0;
    c->k.row_stride = 1;
    c->k.row_block = 1;
    c->k.div_x = cfg->uv_div_x;
    c->k.div_y = cfg->uv_div_y;
    for (int i = 0; i < 3; ++i) {
        c->k.cam[i] = cfg->cam[i];
        c->k.centre[i] = cfg->centre[i];
    }
    c->k.radius = cfg->radius;

    auto bail = [&](cudaError_t err, const char *what) {
        int rc = cuda_fail(nullptr, err, what);
        ptb_destroy(c);
        return rc;
    };
    if (cfg->flags & PTB_CFG_EXTERNAL_STREAM) {
        c->stream = (cudaStream_t)cfg->stream;
    } else {
        if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess)
            return bail(e, "cudaStreamCreate");
        c->own_stream = true;
    }
    if ((e = cudaEventCreate(&c->ev0)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaEventCreate(&c->ev1)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate(copy)");
    if ((e = cudaEventCreateWithFlags(&c->ev_img_ready, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaEventCreateWithFlags(&c->ev_copy_done, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = ptb::image_tables_init()) != cudaSuccess) return bail(e, "image_tables_init");
    if ((e = ptb::prepare_terms(&c->k, c->stream)) != cudaSuccess) return bail(e, "prepare_terms");
    derive_fast_flags(c);

    const size_t n = n_pixels(c);
    // ≙ CreateResources (Renderer.cs:120-179): the W x H RGBA8 storage image.
    if ((e = cudaMalloc(&c->buf.rgba8, n * 4)) != cudaSuccess) return bail(e, "cudaMalloc(rgba8)");
    if ((e = cudaMemsetAsync(c->buf.rgba8, 0, n * 4, c->stream)) != cudaSuccess) return bail(e, "cudaMemset(rgba8)");
    if (c->k.fast && c->k.cm_on && (cfg->width & 3u) == 0) {
        // which quads of every row can skip the intersection altogether: evaluated once, the parameters never change
        if ((e = cudaMalloc(&c->buf.row_span, (size_t)cfg->height * sizeof(uint2))) != cudaSuccess) return bail(e, "cudaMalloc(row_span)");
        if ((e = ptb::launch_row_span(c->k, c->buf.row_span, c->stream)) != cudaSuccess) return bail(e, "row_span_kernel");
    }
    if (cfg->flags & PTB_CFG_KEEP_F32)
        if ((e = cudaMalloc(&c->buf.f32, n * sizeof(float4))) != cudaSuccess) return bail(e, "cudaMalloc(f32)");
    if (cfg->flags & PTB_CFG_KEEP_HIT_IDS)
        if ((e = cudaMalloc(&c->buf.hit, n * sizeof(int32_t))) != cudaSuccess) return bail(e, "cudaMalloc(hit)");
    *out_ctx = c;
    return PTB_OK;
}

int ptb_destroy(ptb_context *ctx)
{
    if (!ctx) return PTB_OK;
    cudaSetDevice(ctx->cfg.device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    ptb::scene_destroy(ctx);
    for (void *p : ctx->ipc_opened) cudaIpcCloseMemHandle(p);
    ctx->ipc_opened.clear();
    if (ctx->peer) {
        ptb::PeerGroup *pg = ctx->peer;
        if (pg->merge_stream) {
            cudaStreamSynchronize(pg->merge_stream);
            cudaStreamDestroy(pg->merge_stream);
        }
        if (pg->partial[1] && pg->partial[1] != ctx->buf.accum_owned) cudaFree(pg->partial[1]);
        if (pg->ev_rendered) cudaEventDestroy(pg->ev_rendered);
        for (int b = 0; b < 2; ++b) {
            if (pg->ev_released[b]) cudaEventDestroy(pg->ev_released[b]);
            if (pg->ev_merged[b]) cudaEventDestroy(pg->ev_merged[b]);
        }
        delete pg;
        ctx->peer = nullptr;
    }
    cudaFree(ctx->slab_total);
    cudaFree(ctx->peer_flags);
    free_buffers(ctx);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->copy_stream) {
        cudaStreamSynchronize(ctx->copy_stream);
        cudaStreamDestroy(ctx->copy_stream);
    }
    for (int i = 0; i < 3; ++i) {
        if (ctx->copy_extra[i]) {
            cudaStreamSynchronize(ctx->copy_extra[i]);
            cudaStreamDestroy(ctx->copy_extra[i]);
        }
        if (ctx->ev_copy_part[i]) cudaEventDestroy(ctx->ev_copy_part[i]);
    }
    if (ctx->ev_img_ready) cudaEventDestroy(ctx->ev_img_ready);
    if (ctx->ev_copy_done) cudaEventDestroy(ctx->ev_copy_done);
    if (ctx->ev_copy_done_alt) cudaEventDestroy(ctx->ev_copy_done_alt);
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return PTB_OK;
}

const char *ptb_last_error(const ptb_context *ctx) { return ctx ? ctx->err.c_str() : g_err.c_str(); }

int ptb_compute_frame(ptb_context *ctx)
{
    PTB_ENTER(ctx);
    if (ctx->copy_pending) PTB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_copy_done, 0));
    if (ptb::scene_resident(ctx)) return ptb::scene_compute_frame(ctx);   // n3-n7: uploaded scene instead of the literal one
    PTB_CUDA(ctx, ptb::launch_csmain_frame(ctx->k, ctx->buf, ctx->stream));
    ctx->launches += 1;
    return PTB_OK;
}

int ptb_compute_frames(ptb_context *ctx, uint32_t n_frames)
{
    PTB_ENTER(ctx);
    if (n_frames == 0) return PTB_OK;
    if (ctx->copy_pending) PTB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_copy_done, 0));
    if (!ptb::scene_resident(ctx)) {
        const cudaError_t e = ptb::launch_csmain_frames(ctx->k, ctx->buf, n_frames, ctx->stream);
        if (e == cudaSuccess) {
            ctx->launches += 1;
            return PTB_OK;
        }
        if (e != cudaErrorNotSupported) return cuda_fail(ctx, e, "launch_csmain_frames");
    }
    for (uint32_t f = 0; f < n_frames; ++f) {      // configurations the persistent kernel does not cover: one launch per frame
        const int rc = ptb_compute_frame(ctx);
        if (rc != PTB_OK) return rc;
    }
    return PTB_OK;
}

int ptb_wait(ptb_context *ctx)
{
    PTB_ENTER(ctx);
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->peer && ctx->peer->merge_stream) {
        PTB_CUDA(ctx, cudaStreamSynchronize(ctx->peer->merge_stream));
        ctx->peer->released_pending[0] = ctx->peer->released_pending[1] = false;
    }
    if (ctx->copy_pending || ctx->copy_pending_alt) {
        PTB_CUDA(ctx, cudaStreamSynchronize(ctx->copy_stream));
        ctx->copy_pending = ctx->copy_pending_alt = false;
    }
    if (ctx->peer_flags) {
        uint32_t err = 0;
        PTB_CUDA(ctx, cudaMemcpy(&err, ctx->peer_flags + 2 * PTB_MAX_PEERS, sizeof err, cudaMemcpyDeviceToHost));
        if (err)
            return fail(ctx, PTB_ERR_STATE, "a cross-GPU rendezvous (ptb_peer_barrier / ptb_peer_frame) timed out waiting for a peer: "
                        "the merges behind it were skipped, the image is incomplete (ptb_accum_reset clears the condition)");
    }
    if (ptb::scene_resident(ctx)) return ptb::scene_wait_check(ctx);
    return PTB_OK;
}

int ptb_read_rgba8_async(ptb_context *ctx, void *host_dst, size_t pitch_bytes)
{
    PTB_ENTER(ctx);
    if (!host_dst) return fail(ctx, PTB_ERR_INVALID, "ptb_read_rgba8_async: host destination is NULL");
    const size_t row = (size_t)ctx->cfg.width * 4;
    if (pitch_bytes == 0) pitch_bytes = row;
    if (pitch_bytes < row) return fail(ctx, PTB_ERR_INVALID, "ptb_read_rgba8_async: pitch %zu < row size %zu", pitch_bytes, row);
    ptb::PeerGroup *pg = ctx->peer;
    const int cur = (pg && pg->bound && ctx->buf.rgba8 == pg->image[1]) ? 1 : 0;
    if (pg && pg->bound && pg->merged_valid[cur]) {
        // the current image was produced on the merge stream (ptb_peer_frame): order the copy after ITS completion, not
        // after the render stream, which is already busy with the next frame
        PTB_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, pg->ev_merged[cur], 0));
    } else {
        PTB_CUDA(ctx, cudaEventRecord(ctx->ev_img_ready, ctx->stream));
        PTB_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_img_ready, 0));
    }
    // The frame goes down in up to 4 row bands on 4 copy streams, so that several copy engines work on one read-back: one
    // engine sustains ~40 GB/s device-to-host on these hosts, which made the 8.3 MB frame of a 1080p step (205 us) the
    // bottleneck of the multi-GPU end-to-end loop; 2 / 4 bands: 179 / 175 us per step next to a 172 us render
    // (experiments/peer_timing.py; PTB200_COPY_SPLIT = 1..4 for A/B runs).  Small images stay on one stream.
    // A context on its own keeps ONE stream: there the copy overlaps the next frame and only its throughput counts — one
    // stream 50.8 GB/s (e2e 812 k Msamples/s on c2), 4 bands 49.4 GB/s (790 k): the bands only add per-copy overheads.
    static const int split_env = [] {
        const char *e = std::getenv("PTB200_COPY_SPLIT");
        const int v = e ? std::atoi(e) : 0;
        return v >= 1 && v <= 4 ? v : 0;
    }();
    const int split = split_env ? split_env : (pg && pg->bound ? 4 : 1);
    const uint32_t H = ctx->cfg.height;
    const int parts = (row * H >= (1u << 20)) ? (int)std::min<uint32_t>((uint32_t)split, H) : 1;
    if (parts > 1) {
        PTB_CUDA(ctx, cudaEventRecord(ctx->ev_img_ready, ctx->copy_stream));       // "the image may be read" as seen by the copy stream
        for (int i = 1; i < parts; ++i) {
            if (!ctx->copy_extra[i - 1]) {
                PTB_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_extra[i - 1], cudaStreamNonBlocking));
                PTB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_copy_part[i - 1], cudaEventDisableTiming));
            }
            const uint32_t r0 = (uint32_t)((uint64_t)H * i / parts), r1 = (uint32_t)((uint64_t)H * (i + 1) / parts);
            PTB_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_extra[i - 1], ctx->ev_img_ready, 0));
            PTB_CUDA(ctx, cudaMemcpy2DAsync(static_cast<char *>(host_dst) + (size_t)r0 * pitch_bytes, pitch_bytes,
                                            reinterpret_cast<const char *>(ctx->buf.rgba8) + (size_t)r0 * row, row, row, r1 - r0,
                                            cudaMemcpyDeviceToHost, ctx->copy_extra[i - 1]));
            PTB_CUDA(ctx, cudaEventRecord(ctx->ev_copy_part[i - 1], ctx->copy_extra[i - 1]));
        }
    }
    PTB_CUDA(ctx, cudaMemcpy2DAsync(host_dst, pitch_bytes, ctx->buf.rgba8, row, row, parts > 1 ? (uint32_t)((uint64_t)H / parts) : H,
                                    cudaMemcpyDeviceToHost, ctx->copy_stream));
    for (int i = 1; i < parts; ++i) PTB_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_copy_part[i - 1], 0));
    PTB_CUDA(ctx, cudaEventRecord(ctx->ev_copy_done, ctx->copy_stream));
    ctx->copy_pending = true;
    return PTB_OK;
}

int ptb_read_rgba8(ptb_context *ctx, void *host_dst, size_t pitch_bytes)
{
    PTB_ENTER(ctx);
    if (ctx->peer && ctx->peer->bound) {
        const int rc = ptb_peer_finish(ctx);
        if (rc != PTB_OK) return rc;
    }
    return read_rows(ctx, host_dst, pitch_bytes, ctx->buf.rgba8, (size_t)ctx->cfg.width * 4, ctx->cfg.height);
}

int ptb_read_f32(ptb_context *ctx, float *host_dst, size_t pitch_bytes)
{
    PTB_ENTER(ctx);
    if (!ctx->buf.f32) return fail(ctx, PTB_ERR_STATE, "ptb_read_f32: context created without PTB_CFG_KEEP_F32");
    return read_rows(ctx, host_dst, pitch_bytes, ctx->buf.f32, (size_t)ctx->cfg.width * 16, ctx->cfg.height);
}

int ptb_read_hit_ids(ptb_context *ctx, int32_t *host_dst, size_t pitch_bytes)
{
    PTB_ENTER(ctx);
    if (!ctx->buf.hit) return fail(ctx, PTB_ERR_STATE, "ptb_read_hit_ids: context created without PTB_CFG_KEEP_HIT_IDS");
    return read_rows(ctx, host_dst, pitch_bytes, ctx->buf.hit, (size_t)ctx->cfg.width * 4, ctx->cfg.height);
}

void *ptb_device_rgba8(ptb_context *ctx) { return ctx ? ctx->buf.rgba8 : nullptr; }
void *ptb_device_f32(ptb_context *ctx) { return ctx ? ctx->buf.f32 : nullptr; }

static int ensure_accum(ptb_context *ctx)
{
    if (ctx->buf.accum) return PTB_OK;
    const size_t bytes = n_pixels(ctx) * sizeof(float4);
    PTB_CUDA(ctx, cudaMalloc(&ctx->buf.accum_owned, bytes));
    ctx->buf.accum = ctx->buf.accum_owned;
    PTB_CUDA(ctx, cudaMemsetAsync(ctx->buf.accum, 0, bytes, ctx->stream));
    ctx->accum_spp = 0;
    return PTB_OK;
}

void *ptb_device_accum(ptb_context *ctx)
{
    if (!ctx) return nullptr;
    if (cudaSetDevice(ctx->cfg.device) != cudaSuccess) return nullptr;
    if (ensure_accum(ctx) != PTB_OK) return nullptr;
    return ctx->buf.accum;
}

int ptb_accum_reset(ptb_context *ctx)
{
    PTB_ENTER(ctx);
    if (ctx->peer && ctx->peer->merge_stream) PTB_CUDA(ctx, cudaStreamSynchronize(ctx->peer->merge_stream));   // merges in flight use these buffers
    int rc = ensure_accum(ctx);
    if (rc != PTB_OK) return rc;
    PTB_CUDA(ctx, cudaMemsetAsync(ctx->buf.accum, 0, n_pixels(ctx) * sizeof(float4), ctx->stream));
    if (ctx->slab_total)
        PTB_CUDA(ctx, cudaMemsetAsync(ctx->slab_total, 0, (size_t)ctx->slab_count * sizeof(float4), ctx->stream));
    if (ctx->peer_flags)     // a timed-out rendezvous is sticky until the progressive state is reset
        PTB_CUDA(ctx, cudaMemsetAsync(ctx->peer_flags + 2 * PTB_MAX_PEERS, 0, sizeof(uint32_t), ctx->stream));
    if (ctx->peer) {
        ptb::PeerGroup *pg = ctx->peer;
        for (int b = 0; b < 2; ++b)
            if (pg->partial[b] && pg->partial[b] != ctx->buf.accum)
                PTB_CUDA(ctx, cudaMemsetAsync(pg->partial[b], 0, n_pixels(ctx) * sizeof(float4), ctx->stream));
        PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        pg->released_pending[0] = pg->released_pending[1] = false;
    }
    ctx->accum_spp = 0;
    return PTB_OK;
}

int ptb_render_samples(ptb_context *ctx, uint32_t first_sample, uint32_t n_samples, int jitter)
{
    PTB_ENTER(ctx);
    if (n_samples == 0) return PTB_OK;
    if ((uint64_t)ctx->accum_spp + n_samples > 0xFFFFFFFFull)
        return fail(ctx, PTB_ERR_INVALID, "ptb_render_samples: spp counter would overflow");
    int rc = ensure_accum(ctx);
    if (rc != PTB_OK) return rc;
    if (ptb::scene_resident(ctx)) {
        rc = ptb::scene_render_samples(ctx, first_sample, n_samples, jitter);
        if (rc == PTB_OK) ctx->accum_spp += n_samples;
        return rc;
    }
    PTB_CUDA(ctx, ptb::launch_csmain_accumulate(ctx->k, ctx->buf.accum, first_sample, n_samples, ctx->cfg.seed,
                                                jitter != 0, ctx->stream));
    ctx->launches += 1;
    ctx->accum_spp += n_samples;
    return PTB_OK;
}

int ptb_render_resolve(ptb_context *ctx, uint32_t first_sample, uint32_t n_samples, int jitter, uint32_t flags)
{
    PTB_ENTER(ctx);
    if (flags & ~(PTB_RESOLVE_SRGB | PTB_RESOLVE_BGRA))
        return fail(ctx, PTB_ERR_INVALID, "ptb_render_resolve: unknown flags 0x%x", flags);
    if (n_samples == 0) return fail(ctx, PTB_ERR_INVALID, "ptb_render_resolve: n_samples is 0");
    // One launch only where the accumulate kernel touches every pixel the resolve would: the literal scene, no row shard,
    // no sRGB table.  Everything else takes the two calls this one stands for.
    const bool fused = !ptb::scene_resident(ctx) && !(flags & PTB_RESOLVE_SRGB) && ctx->k.row_offset == 0 &&
                       ctx->k.row_stride == 1;
    if (!fused) {
        const int rc = ptb_render_samples(ctx, first_sample, n_samples, jitter);
        return rc != PTB_OK ? rc : ptb_resolve(ctx, flags);
    }
    if ((uint64_t)ctx->accum_spp + n_samples > 0xFFFFFFFFull)
        return fail(ctx, PTB_ERR_INVALID, "ptb_render_resolve: spp counter would overflow");
    int rc = ensure_accum(ctx);
    if (rc != PTB_OK) return rc;
    if (!ctx->buf.rgba8_alt) {
        PTB_CUDA(ctx, cudaMalloc(&ctx->buf.rgba8_alt, n_pixels(ctx) * sizeof(uint32_t)));
        PTB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_copy_done_alt, cudaEventDisableTiming));
    }
    // the image this launch writes is the one that was current two calls ago: a read-back of it may still be running
    if (ctx->copy_pending_alt) PTB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_copy_done_alt, 0));
    PTB_CUDA(ctx, ptb::launch_csmain_accumulate(ctx->k, ctx->buf.accum, first_sample, n_samples, ctx->cfg.seed, jitter != 0,
                                                ctx->stream, ctx->buf.rgba8_alt, (float)(ctx->accum_spp + n_samples),
                                                (flags & PTB_RESOLVE_BGRA) != 0));
    std::swap(ctx->buf.rgba8, ctx->buf.rgba8_alt);
    std::swap(ctx->ev_copy_done, ctx->ev_copy_done_alt);
    ctx->copy_pending_alt = ctx->copy_pending;      // a read-back of the image that was current until now
    ctx->copy_pending = false;                      // the new current image: its old read-back is ordered before this launch
    ctx->launches += 1;
    ctx->accum_spp += n_samples;
    return PTB_OK;
}

int ptb_resolve(ptb_context *ctx, uint32_t flags)
{
    PTB_ENTER(ctx);
    if (flags & ~(PTB_RESOLVE_SRGB | PTB_RESOLVE_BGRA)) return fail(ctx, PTB_ERR_INVALID, "ptb_resolve: unknown flags 0x%x", flags);
    if (!ctx->buf.accum || ctx->accum_spp == 0)
        return fail(ctx, PTB_ERR_STATE, "ptb_resolve: nothing accumulated (0 spp)");
    if (ctx->copy_pending) PTB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_copy_done, 0));
    PTB_CUDA(ctx, ptb::launch_resolve(ctx->buf.accum, ctx->buf.rgba8, (uint32_t)n_pixels(ctx), ctx->accum_spp, flags,
                                      ctx->stream));
    ctx->launches += 1;
    return PTB_OK;
}

uint32_t ptb_accum_spp(const ptb_context *ctx) { return ctx ? ctx->accum_spp : 0; }

int ptb_read_accum(ptb_context *ctx, float *host_dst, size_t pitch_bytes)
{
    PTB_ENTER(ctx);
    if (!ctx->buf.accum) return fail(ctx, PTB_ERR_STATE, "ptb_read_accum: no accumulation buffer yet");
    return read_rows(ctx, host_dst, pitch_bytes, ctx->buf.accum, (size_t)ctx->cfg.width * 16, ctx->cfg.height);
}

int ptb_write_accum(ptb_context *ctx, const float *host_src, size_t pitch_bytes, uint32_t spp_total)
{
    PTB_ENTER(ctx);
    if (!host_src) return fail(ctx, PTB_ERR_INVALID, "ptb_write_accum: host_src is NULL");
    const size_t row = (size_t)ctx->cfg.width * 16;
    if (pitch_bytes == 0) pitch_bytes = row;
    if (pitch_bytes < row) return fail(ctx, PTB_ERR_INVALID, "ptb_write_accum: pitch %zu < row size %zu", pitch_bytes, row);
    int rc = ensure_accum(ctx);
    if (rc != PTB_OK) return rc;
    PTB_CUDA(ctx, cudaMemcpy2DAsync(ctx->buf.accum, row, host_src, pitch_bytes, row, ctx->cfg.height,
                                    cudaMemcpyHostToDevice, ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->accum_spp = spp_total;
    return PTB_OK;
}

int ptb_set_accum_buffer(ptb_context *ctx, void *device_ptr)
{
    PTB_ENTER(ctx);
    if (device_ptr) {
        if ((uintptr_t)device_ptr & 15u) return fail(ctx, PTB_ERR_INVALID, "ptb_set_accum_buffer: pointer not 16 B aligned");
        cudaPointerAttributes attr{};
        PTB_CUDA(ctx, cudaPointerGetAttributes(&attr, device_ptr));
        if (attr.type != cudaMemoryTypeDevice && attr.type != cudaMemoryTypeManaged)
            return fail(ctx, PTB_ERR_INVALID, "ptb_set_accum_buffer: not a device pointer");
        ctx->buf.accum = static_cast<float4 *>(device_ptr);
    } else {
        ctx->buf.accum = ctx->buf.accum_owned;   // may be NULL: allocated lazily again
    }
    ctx->accum_spp = 0;
    return PTB_OK;
}

int ptb_set_row_shard_bands(ptb_context *ctx, uint32_t row_offset, uint32_t row_stride, uint32_t rows_per_band)
{
    PTB_ENTER(ctx);
    if (row_stride == 0 || row_offset >= row_stride || rows_per_band == 0)
        return fail(ctx, PTB_ERR_INVALID, "ptb_set_row_shard: need row_stride >= 1, row_offset < row_stride and rows_per_band >= 1 "
                    "(got %u, %u, %u)", row_offset, row_stride, rows_per_band);
    ctx->k.row_offset = row_offset;
    ctx->k.row_stride = row_stride;
    ctx->k.row_block = rows_per_band;
    return PTB_OK;
}

int ptb_set_row_shard(ptb_context *ctx, uint32_t row_offset, uint32_t row_stride)
{
    return ptb_set_row_shard_bands(ctx, row_offset, row_stride, 1u);
}

int ptb_set_accum_spp(ptb_context *ctx, uint32_t spp_total)
{
    PTB_ENTER(ctx);
    ctx->accum_spp = spp_total;
    return PTB_OK;
}

static int ensure_peer_flags(ptb_context *ctx)
{
    if (ctx->peer_flags) return PTB_OK;
    const size_t bytes = (2 * PTB_MAX_PEERS + 1) * sizeof(uint32_t);
    PTB_CUDA(ctx, cudaMalloc(&ctx->peer_flags, bytes));
    PTB_CUDA(ctx, cudaMemset(ctx->peer_flags, 0, bytes));
    return PTB_OK;
}

void *ptb_device_peer_flags(ptb_context *ctx)
{
    if (!ctx || cudaSetDevice(ctx->cfg.device) != cudaSuccess || ensure_peer_flags(ctx) != PTB_OK) return nullptr;
    return ctx->peer_flags;
}

int ptb_peer_barrier(ptb_context *ctx, uint32_t n_peers, void *const *peer_flags, uint32_t self_rank, uint32_t phase,
                     uint32_t epoch)
{
    PTB_ENTER(ctx);
    if (n_peers == 0 || n_peers > PTB_MAX_PEERS || !peer_flags || self_rank >= n_peers || phase > 1)
        return fail(ctx, PTB_ERR_INVALID, "ptb_peer_barrier: bad argument");
    for (uint32_t q = 0; q < n_peers; ++q)
        if (!peer_flags[q]) return fail(ctx, PTB_ERR_INVALID, "ptb_peer_barrier: peer %u pointer is NULL", q);
    int rc = ensure_peer_flags(ctx);
    if (rc != PTB_OK) return rc;
    if (peer_flags[self_rank] != ctx->peer_flags)
        return fail(ctx, PTB_ERR_INVALID, "ptb_peer_barrier: peer_flags[self_rank] must be this context's own flag block");
    PTB_CUDA(ctx, ptb::launch_peer_barrier(peer_flags, n_peers, self_rank, phase, epoch, ctx->peer_flags + 2 * PTB_MAX_PEERS,
                                           ctx->stream));
    ctx->launches += 1;
    return PTB_OK;
}

int ptb_accum_clear(ptb_context *ctx)
{
    PTB_ENTER(ctx);
    int rc = ensure_accum(ctx);
    if (rc != PTB_OK) return rc;
    // a kernel rather than a memset: it does nothing when the rendezvous before it timed out (the partials then survive)
    PTB_CUDA(ctx, ptb::launch_accum_clear(ctx->buf.accum, (uint32_t)n_pixels(ctx),
                                          ctx->peer_flags ? ctx->peer_flags + 2 * PTB_MAX_PEERS : nullptr, ctx->stream));
    ctx->launches += 1;
    return PTB_OK;
}

int ptb_ipc_export(ptb_context *ctx, int which, void *handle_out)
{
    PTB_ENTER(ctx);
    if (!handle_out || which < 0 || which > 4) return fail(ctx, PTB_ERR_INVALID, "ptb_ipc_export: bad argument");
    if (which >= 3 && !ctx->peer) return fail(ctx, PTB_ERR_STATE, "ptb_ipc_export: which = 3 / 4 need ptb_peer_group_create");
    static_assert(sizeof(cudaIpcMemHandle_t) == PTB_IPC_HANDLE_BYTES, "IPC handle size");
    void *ptr = nullptr;
    if (which == 0) {
        int rc = ensure_accum(ctx);
        if (rc != PTB_OK) return rc;
        if (ctx->buf.accum != ctx->buf.accum_owned)
            return fail(ctx, PTB_ERR_STATE, "ptb_ipc_export: the accumulation buffer is caller-owned (ptb_set_accum_buffer)");
        ptr = ctx->buf.accum_owned;
    } else if (which == 1) {
        ptr = ctx->peer ? ctx->peer->image[0] : ctx->buf.rgba8;
    } else if (which == 2) {
        int rc = ensure_peer_flags(ctx);
        if (rc != PTB_OK) return rc;
        ptr = ctx->peer_flags;
    } else if (which == 3) {
        ptr = ctx->peer->partial[1];
    } else {
        ptr = ctx->peer->image[1];
    }
    cudaIpcMemHandle_t h;
    PTB_CUDA(ctx, cudaIpcGetMemHandle(&h, ptr));
    std::memcpy(handle_out, &h, sizeof h);
    return PTB_OK;
}

int ptb_ipc_open(ptb_context *ctx, const void *handle, void **device_ptr_out)
{
    PTB_ENTER(ctx);
    if (!handle || !device_ptr_out) return fail(ctx, PTB_ERR_INVALID, "ptb_ipc_open: NULL argument");
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof h);
    void *p = nullptr;
    PTB_CUDA(ctx, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    ctx->ipc_opened.push_back(p);
    *device_ptr_out = p;
    return PTB_OK;
}

int ptb_peer_merge(ptb_context *ctx, uint32_t n_peers, void *const *peer_accum, void *dst_image, uint32_t slab_first,
                   uint32_t slab_count, uint32_t spp_total, uint32_t flags)
{
    PTB_ENTER(ctx);
    if (n_peers == 0 || n_peers > PTB_MAX_PEERS || !peer_accum || !dst_image)
        return fail(ctx, PTB_ERR_INVALID, "ptb_peer_merge: need 1..%d peers and a destination image", PTB_MAX_PEERS);
    if (flags & ~(PTB_RESOLVE_SRGB | PTB_RESOLVE_BGRA | PTB_MERGE_BACKGROUND))
        return fail(ctx, PTB_ERR_INVALID, "ptb_peer_merge: unknown flags 0x%x", flags);
    if (spp_total == 0) return fail(ctx, PTB_ERR_STATE, "ptb_peer_merge: 0 spp");
    if ((uint64_t)slab_first + slab_count > n_pixels(ctx)) return fail(ctx, PTB_ERR_INVALID, "ptb_peer_merge: slab beyond the image");
    for (uint32_t q = 0; q < n_peers; ++q)
        if (!peer_accum[q] || ((uintptr_t)peer_accum[q] & 15u)) return fail(ctx, PTB_ERR_INVALID, "ptb_peer_merge: peer %u pointer", q);
    if (slab_count == 0) return PTB_OK;
    if (!ctx->slab_total || ctx->slab_first != slab_first || ctx->slab_count != slab_count) {
        cudaFree(ctx->slab_total);
        ctx->slab_total = nullptr;
        PTB_CUDA(ctx, cudaMalloc(&ctx->slab_total, (size_t)slab_count * sizeof(float4)));
        PTB_CUDA(ctx, cudaMemsetAsync(ctx->slab_total, 0, (size_t)slab_count * sizeof(float4), ctx->stream));
        ctx->slab_first = slab_first;
        ctx->slab_count = slab_count;
    }
    PTB_CUDA(ctx, ptb::launch_peer_merge(peer_accum, n_peers, ctx->slab_total, static_cast<uint32_t *>(dst_image), slab_first,
                                         slab_count, spp_total, flags, ctx->stream,
                                         ctx->peer_flags ? ctx->peer_flags + 2 * PTB_MAX_PEERS : nullptr));
    ctx->launches += 1;
    return PTB_OK;
}

// ---- native multi-GPU frame loop ---------------------------------------------------------------------------------
static void slab_of(uint32_t rank, uint32_t world, uint64_t n, uint32_t *first, uint32_t *count)
{
    const uint64_t base = n / world, extra = n % world;        // the split of pathtracing_b200/dist.py sample_range
    *count = (uint32_t)(base + (rank < extra ? 1u : 0u));
    *first = (uint32_t)(rank * base + (rank < extra ? rank : extra));
}

int ptb_peer_group_create(ptb_context *ctx)
{
    PTB_ENTER(ctx);
    if (ctx->peer) return PTB_OK;
    int rc = ensure_accum(ctx);
    if (rc != PTB_OK) return rc;
    if (ctx->buf.accum != ctx->buf.accum_owned)
        return fail(ctx, PTB_ERR_STATE, "ptb_peer_group_create: the accumulation buffer is caller-owned (ptb_set_accum_buffer)");
    rc = ensure_peer_flags(ctx);
    if (rc != PTB_OK) return rc;
    if (!ctx->buf.rgba8_alt) {
        PTB_CUDA(ctx, cudaMalloc(&ctx->buf.rgba8_alt, n_pixels(ctx) * sizeof(uint32_t)));
        PTB_CUDA(ctx, cudaMemsetAsync(ctx->buf.rgba8_alt, 0, n_pixels(ctx) * sizeof(uint32_t), ctx->stream));
        PTB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_copy_done_alt, cudaEventDisableTiming));
    }
    ptb::PeerGroup *pg = new (std::nothrow) ptb::PeerGroup();
    if (!pg) return fail(ctx, PTB_ERR_NOMEM, "ptb_peer_group_create: host allocation failed");
    ctx->peer = pg;                                  // from here on ptb_destroy releases whatever was created
    int lo = 0, hi = 0;
    PTB_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&lo, &hi));
    PTB_CUDA(ctx, cudaStreamCreateWithPriority(&pg->merge_stream, cudaStreamNonBlocking, hi));
    pg->partial[0] = ctx->buf.accum_owned;
    PTB_CUDA(ctx, cudaMalloc(&pg->partial[1], n_pixels(ctx) * sizeof(float4)));
    PTB_CUDA(ctx, cudaMemsetAsync(pg->partial[1], 0, n_pixels(ctx) * sizeof(float4), ctx->stream));
    pg->image[0] = ctx->buf.rgba8;
    pg->image[1] = ctx->buf.rgba8_alt;
    PTB_CUDA(ctx, cudaEventCreateWithFlags(&pg->ev_rendered, cudaEventDisableTiming));
    for (int b = 0; b < 2; ++b) {
        PTB_CUDA(ctx, cudaEventCreateWithFlags(&pg->ev_released[b], cudaEventDisableTiming));
        PTB_CUDA(ctx, cudaEventCreateWithFlags(&pg->ev_merged[b], cudaEventDisableTiming));
    }
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PTB_OK;
}

void *ptb_peer_group_local(ptb_context *ctx, int which)
{
    if (!ctx || !ctx->peer) return nullptr;
    switch (which) {
    case 0: return ctx->peer->partial[0];
    case 1: return ctx->peer->image[0];
    case 2: return ctx->peer_flags;
    case 3: return ctx->peer->partial[1];
    case 4: return ctx->peer->image[1];
    default: return nullptr;
    }
}

int ptb_peer_group_bind(ptb_context *ctx, uint32_t n_peers, uint32_t self_rank, uint32_t dst_rank, void *const *partial0,
                        void *const *partial1, void *const *flags, void *dst_image0, void *dst_image1)
{
    PTB_ENTER(ctx);
    if (!ctx->peer) return fail(ctx, PTB_ERR_STATE, "ptb_peer_group_bind: call ptb_peer_group_create first");
    if (n_peers == 0 || n_peers > PTB_MAX_PEERS || self_rank >= n_peers || dst_rank >= n_peers || !partial0 || !partial1 || !flags ||
        !dst_image0 || !dst_image1)
        return fail(ctx, PTB_ERR_INVALID, "ptb_peer_group_bind: bad argument");
    ptb::PeerGroup *pg = ctx->peer;
    for (uint32_t q = 0; q < n_peers; ++q) {
        if (!partial0[q] || !partial1[q] || !flags[q] || ((uintptr_t)partial0[q] & 15u) || ((uintptr_t)partial1[q] & 15u))
            return fail(ctx, PTB_ERR_INVALID, "ptb_peer_group_bind: peer %u pointer is NULL or misaligned", q);
        // peers of the SAME process on other devices (one host process driving several GPUs, as the reference's C# host
        // would): plain device pointers, which need peer access switched on; IPC mappings already carry it
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, partial0[q]) == cudaSuccess && at.type == cudaMemoryTypeDevice && at.device != ctx->cfg.device) {
            int can = 0;
            PTB_CUDA(ctx, cudaDeviceCanAccessPeer(&can, ctx->cfg.device, at.device));
            if (!can) return fail(ctx, PTB_ERR_UNSUPPORTED, "ptb_peer_group_bind: device %d cannot access peer %u's device %d", ctx->cfg.device, q, at.device);
            const cudaError_t pe = cudaDeviceEnablePeerAccess(at.device, 0);
            if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) PTB_CUDA(ctx, pe);
            (void)cudaGetLastError();
        } else {
            (void)cudaGetLastError();
        }
        pg->peer_partial[0][q] = partial0[q];
        pg->peer_partial[1][q] = partial1[q];
        pg->peer_flags[q] = flags[q];
    }
    if (partial0[self_rank] != pg->partial[0] || partial1[self_rank] != pg->partial[1] || flags[self_rank] != ctx->peer_flags)
        return fail(ctx, PTB_ERR_INVALID, "ptb_peer_group_bind: entry self_rank of every array must be this context's own buffer");
    if (dst_rank == self_rank && (dst_image0 != pg->image[0] || dst_image1 != pg->image[1]))
        return fail(ctx, PTB_ERR_INVALID, "ptb_peer_group_bind: the destination images of the destination rank must be its own");
    pg->dst_image[0] = dst_image0;
    pg->dst_image[1] = dst_image1;
    pg->n_peers = n_peers;
    pg->self = self_rank;
    pg->dst = dst_rank;
    uint32_t first = 0, count = 0;
    slab_of(self_rank, n_peers, n_pixels(ctx), &first, &count);
    if (!ctx->slab_total || ctx->slab_first != first || ctx->slab_count != count) {
        cudaFree(ctx->slab_total);
        ctx->slab_total = nullptr;
        PTB_CUDA(ctx, cudaMalloc(&ctx->slab_total, (size_t)(count ? count : 1) * sizeof(float4)));
        ctx->slab_first = first;
        ctx->slab_count = count;
    }
    PTB_CUDA(ctx, cudaMemsetAsync(ctx->slab_total, 0, (size_t)(count ? count : 1) * sizeof(float4), ctx->stream));
    // a (re-)bind restarts the rendezvous epochs: own flag block and error word back to zero.  Every rank binds, then the
    // host synchronises the ranks once (dist.py: barrier) before the first frame, so no signal can precede this reset.
    PTB_CUDA(ctx, cudaStreamSynchronize(pg->merge_stream));
    PTB_CUDA(ctx, cudaMemsetAsync(ctx->peer_flags, 0, (2 * PTB_MAX_PEERS + 1) * sizeof(uint32_t), ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    pg->epoch[0] = pg->epoch[1] = 0;
    pg->frame = 0;
    pg->released_pending[0] = pg->released_pending[1] = false;
    pg->merged_valid[0] = pg->merged_valid[1] = false;
    pg->bound = true;
    return PTB_OK;
}

int ptb_peer_frame(ptb_context *ctx, uint32_t first_sample, uint32_t n_samples, int jitter, uint32_t spp_total, uint32_t flags)
{
    PTB_ENTER(ctx);
    ptb::PeerGroup *pg = ctx->peer;
    if (!pg || !pg->bound) return fail(ctx, PTB_ERR_STATE, "ptb_peer_frame: no peer group bound (ptb_peer_group_bind)");
    if (flags & ~(PTB_RESOLVE_SRGB | PTB_RESOLVE_BGRA | PTB_PEER_RENDER_ONLY))
        return fail(ctx, PTB_ERR_INVALID, "ptb_peer_frame: unknown flags 0x%x", flags);
    const bool render_only = (flags & PTB_PEER_RENDER_ONLY) != 0;
    flags &= ~PTB_PEER_RENDER_ONLY;
    if (spp_total == 0 && !render_only) return fail(ctx, PTB_ERR_INVALID, "ptb_peer_frame: spp_total is 0");
    const int b = (int)(pg->frame & 1u);
    uint32_t *const err = ctx->peer_flags + 2 * PTB_MAX_PEERS;
    // ---- render stream: this rank's samples into partial[b], once the merge of frame k-2 has released it
    if (pg->released_pending[b]) {
        PTB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, pg->ev_released[b], 0));
        pg->released_pending[b] = false;
    }
    ctx->buf.accum = pg->partial[b];
    if (n_samples) {
        const uint32_t spp_before = ctx->accum_spp;
        ctx->accum_spp = 0;                               // the partial holds this frame's samples only
        const int rc = ptb_render_samples(ctx, first_sample, n_samples, jitter);
        ctx->accum_spp = spp_before;
        if (rc != PTB_OK) return rc;
    }
    if (render_only) return PTB_OK;                       // more samples of this frame follow; a later call merges
    PTB_CUDA(ctx, cudaEventRecord(pg->ev_rendered, ctx->stream));
    // ---- merge stream: rendezvous, fused merge + resolve of this rank's slab into dst's image[b], rendezvous, clear
    cudaStream_t ms = pg->merge_stream;
    PTB_CUDA(ctx, cudaStreamWaitEvent(ms, pg->ev_rendered, 0));
    if (pg->self == pg->dst) {
        // image[b] is about to be overwritten by every rank: a read-back of it (frame k-2) must have finished first
        const bool is_cur = ctx->buf.rgba8 == pg->image[b];
        const bool pending = is_cur ? ctx->copy_pending : ctx->copy_pending_alt;
        if (pending) PTB_CUDA(ctx, cudaStreamWaitEvent(ms, is_cur ? ctx->ev_copy_done : ctx->ev_copy_done_alt, 0));
    }
    PTB_CUDA(ctx, ptb::launch_peer_barrier(pg->peer_flags, pg->n_peers, pg->self, 0, ++pg->epoch[0], err, ms));
    PTB_CUDA(ctx, ptb::launch_peer_merge(pg->peer_partial[b], pg->n_peers, ctx->slab_total, static_cast<uint32_t *>(pg->dst_image[b]),
                                         ctx->slab_first, ctx->slab_count, spp_total,
                                         flags | (pg->n_peers > 1 ? PTB_MERGE_BACKGROUND : 0u), ms, err));
    PTB_CUDA(ctx, ptb::launch_peer_barrier(pg->peer_flags, pg->n_peers, pg->self, 1, ++pg->epoch[1], err, ms));
    PTB_CUDA(ctx, cudaEventRecord(pg->ev_merged[b], ms));
    pg->merged_valid[b] = true;
    PTB_CUDA(ctx, ptb::launch_accum_clear(pg->partial[b], (uint32_t)n_pixels(ctx), err, ms));
    PTB_CUDA(ctx, cudaEventRecord(pg->ev_released[b], ms));
    pg->released_pending[b] = true;
    ctx->launches += 4;
    // image[b] becomes the current image (ptb_read_rgba8*, ptb_device_rgba8), with its own read-back bookkeeping
    if (ctx->buf.rgba8 != pg->image[b]) {
        std::swap(ctx->buf.rgba8, ctx->buf.rgba8_alt);
        std::swap(ctx->ev_copy_done, ctx->ev_copy_done_alt);
        std::swap(ctx->copy_pending, ctx->copy_pending_alt);
    }
    if (pg->self == pg->dst) ctx->copy_pending = false;   // its last read-back is ordered before the merge that refills it
    pg->frame += 1;
    return PTB_OK;
}

int ptb_peer_finish(ptb_context *ctx)
{
    PTB_ENTER(ctx);
    ptb::PeerGroup *pg = ctx->peer;
    if (!pg) return fail(ctx, PTB_ERR_STATE, "ptb_peer_finish: no peer group");
    for (int b = 0; b < 2; ++b)
        if (pg->released_pending[b]) PTB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, pg->ev_released[b], 0));
    return PTB_OK;
}

int ptb_peer_read_total(ptb_context *ctx, float *host_dst, uint32_t *slab_first, uint32_t *slab_count)
{
    PTB_ENTER(ctx);
    if (!ctx->slab_total) return fail(ctx, PTB_ERR_STATE, "ptb_peer_read_total: no merge has run on this context");
    if (slab_first) *slab_first = ctx->slab_first;
    if (slab_count) *slab_count = ctx->slab_count;
    if (!host_dst) return PTB_OK;
    if (ctx->peer && ctx->peer->merge_stream) PTB_CUDA(ctx, cudaStreamSynchronize(ctx->peer->merge_stream));
    PTB_CUDA(ctx, cudaMemcpyAsync(host_dst, ctx->slab_total, (size_t)ctx->slab_count * sizeof(float4), cudaMemcpyDeviceToHost,
                                  ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PTB_OK;
}

int ptb_present(ptb_context *ctx, uint32_t dst_w, uint32_t dst_h, uint32_t flags)
{
    PTB_ENTER(ctx);
    if (dst_w == 0 || dst_h == 0) return fail(ctx, PTB_ERR_INVALID, "ptb_present: empty target");
    if (flags & ~PTB_RESOLVE_SRGB) return fail(ctx, PTB_ERR_INVALID, "ptb_present: unknown flags 0x%x", flags);
    if ((uint64_t)dst_w * dst_h > (1ull << 30)) return fail(ctx, PTB_ERR_UNSUPPORTED, "ptb_present: target too large");
    const size_t need = (size_t)dst_w * dst_h * 4;
    if (need > ctx->buf.present_cap) {
        cudaFree(ctx->buf.present);
        ctx->buf.present = nullptr;
        ctx->buf.present_cap = 0;
        PTB_CUDA(ctx, cudaMalloc(&ctx->buf.present, need));
        ctx->buf.present_cap = need;
    }
    ctx->buf.present_w = dst_w;
    ctx->buf.present_h = dst_h;
    PTB_CUDA(ctx, ptb::launch_present(ctx->buf.rgba8, ctx->cfg.width, ctx->cfg.height, ctx->buf.present, dst_w, dst_h,
                                      flags, ctx->stream));
    ctx->launches += 1;
    return PTB_OK;
}

int ptb_read_present(ptb_context *ctx, void *host_dst, size_t pitch_bytes)
{
    PTB_ENTER(ctx);
    if (!ctx->buf.present || ctx->buf.present_w == 0) return fail(ctx, PTB_ERR_STATE, "ptb_read_present: ptb_present not called");
    return read_rows(ctx, host_dst, pitch_bytes, ctx->buf.present, (size_t)ctx->buf.present_w * 4, ctx->buf.present_h);
}

int ptb_timer_begin(ptb_context *ctx)
{
    PTB_ENTER(ctx);
    PTB_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    ctx->timer_open = true;
    return PTB_OK;
}

int ptb_timer_end(ptb_context *ctx, float *elapsed_ms)
{
    PTB_ENTER(ctx);
    if (!elapsed_ms) return fail(ctx, PTB_ERR_INVALID, "ptb_timer_end: elapsed_ms is NULL");
    if (!ctx->timer_open) return fail(ctx, PTB_ERR_STATE, "ptb_timer_end without ptb_timer_begin");
    PTB_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    PTB_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
    PTB_CUDA(ctx, cudaEventElapsedTime(elapsed_ms, ctx->ev0, ctx->ev1));
    ctx->timer_open = false;
    return PTB_OK;
}

uint64_t ptb_launch_count(const ptb_context *ctx) { return ctx ? ctx->launches : 0; }

int ptb_kernel_mode(const ptb_context *ctx)
{
    if (!ctx) return fail(nullptr, PTB_ERR_INVALID, "ptb_kernel_mode: ctx is NULL");
    if (ctx->k.spirv) return 3;
    return ctx->k.fast ? (ctx->k.on_axis ? 2 : 1) : 0;
}

int ptb_selftest_exact_math(ptb_context *ctx, int which, float divisor, uint64_t *mismatches, uint64_t *tested)
{
    PTB_ENTER(ctx);
    if (which < 0 || which > 5 || !mismatches || !tested)
        return fail(ctx, PTB_ERR_INVALID, "ptb_selftest_exact_math: bad argument");
    if (which == 4) {
        unsigned long long bad4 = 0, n4 = 0;
        PTB_CUDA(ctx, ptb::selftest_row_span(ctx->k, ctx->buf.row_span, &bad4, &n4, ctx->stream));
        ctx->launches += 1;
        *mismatches = bad4;
        *tested = n4;
        return PTB_OK;
    }
    if (which == 3) {
        unsigned long long bad3 = 0, n3 = 0;
        PTB_CUDA(ctx, ptb::selftest_certain_miss(ctx->k, &bad3, &n3, ctx->stream));
        ctx->launches += 1;
        *mismatches = bad3;
        *tested = n3;
        return PTB_OK;
    }
    if ((which == 2 || which == 5) && !(std::fabs(divisor) > 0.0f && std::isfinite(divisor)))
        return fail(ctx, PTB_ERR_INVALID, "ptb_selftest_exact_math: divisor must be finite and non-zero");
    unsigned long long bad = 0, n = 0;
    PTB_CUDA(ctx, ptb::selftest_exact(which, divisor, &bad, &n, ctx->stream));
    ctx->launches += 1;
    *mismatches = bad;
    *tested = n;
    return PTB_OK;
}

}  // extern "C"
