// ptb_internal.h — context layout and kernel-launcher declarations shared by the
// translation units of libptb200.so.  Not part of the public ABI (include/ptb200.h).
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include <cuda_runtime.h>

#include "../../include/ptb200.h"
#include "row_shard.h"

struct ptb_context;

namespace ptb {

// Launch-uniform parameters of the CSMain-equivalent kernels (passed by value as a kernel
// argument -> constant bank).  Literals of Test.hlsl:6,8,12,13 made runtime parameters, plus
// the launch-uniform subexpressions of Test.hlsl:16,19, which are evaluated ONCE ON THE DEVICE
// (prepare_terms kernel, same *_rn intrinsics as the per-pixel code) when the context is created.
struct FrameK {
    uint32_t width, height;
    uint32_t row_offset, row_stride;   // this context renders row BANDS row_offset + i*row_stride (default 0,1 = all)
    uint32_t row_block;                // rows per band (default 1: single interleaved rows)
    float div_x, div_y;
    float cam[3];
    float centre[3];
    float radius;
    // derived on the device:
    float inv_div_x, inv_div_y;   // RN(1/div)
    float dl_x, dl_y;             // RN(1/div - RN(1/div)): low part of the reciprocal (host, from binary64)
    uint32_t div2;                // 1: the 2-instruction constant division is verified for both divisors
    uint32_t one;                 // always 1 (csmain_accumulate: integer adds issued as multiply-adds on the FMA pipe)
    float oc[3];                  // :16 cam - centre
    float c;                      // :19 dot(oc,oc) - r*r
    float c4;                     // 4*c (exact scaling)
    float neg2ocz;                // -(2*oc.z) (exact), used when on_axis
    float nzero;                  // -0.0f at run time: the addend of the un-contractable packed multiply (ptb_math.cuh)
    // launch-uniform validity of the short exact sequences (api.cu: derive_fast_flags)
    uint32_t fast;                // 1: uv division / sqrt / rcp domains verified for these parameters
    uint32_t on_axis;             // 1: oc.x == oc.y == 0 and oc.z is a moderate normal number
    uint32_t spirv;               // 1: PTB_CFG_SPIRV_ARITH — the as-compiled operation list of Compiled/Ray/Test.spirv
    // conservative per-pixel "every sample of this pixel misses" classification (kernels_csmain.cu: certain_miss)
    uint32_t cm_on;               // 1: enabled (fast mode, finite terms)
    float cm_sx, cm_sy;           // 2/div: uv per pixel
    float cm_K;                   // |c| + |oc|^2: scale of the discriminant's terms
};

struct Buffers {
    uint32_t *rgba8 = nullptr;   // W*H packed R8G8B8A8 (Renderer.cs:124): the CURRENT image
    uint32_t *rgba8_alt = nullptr; // second image, allocated by the first ptb_render_resolve (which writes it and swaps)
    float4 *f32 = nullptr;       // W*H float4 (Test.hlsl:1), optional
    int32_t *hit = nullptr;      // W*H hit ids, optional
    uint2 *row_span = nullptr;   // per image row: [first, end) quad columns that are not certain misses (fast modes, width % 4 == 0)
    float4 *accum = nullptr;     // W*H float4 accumulation (n2)
    float4 *accum_owned = nullptr;
    uint32_t *present = nullptr; // dst_w*dst_h packed B8G8R8A8
    uint32_t present_w = 0, present_h = 0;
    size_t present_cap = 0;
};

// kernels_csmain.cu
cudaError_t prepare_terms(FrameK *k_host, cudaStream_t s);   // fills the derived fields (device-evaluated)
cudaError_t launch_csmain_frame(const FrameK &k, const Buffers &b, cudaStream_t s);
// n_frames literal frames in ONE persistent launch (one resident wave of blocks looping over 1024-pixel row chunks of
// every frame): RGBA8 only, width % 4 == 0, no row shard.  Returns cudaErrorNotSupported when the configuration needs
// the per-frame kernels.
cudaError_t launch_csmain_frames(const FrameK &k, const Buffers &b, uint32_t n_frames, cudaStream_t s);
cudaError_t launch_row_span(const FrameK &k, uint2 *row_span, cudaStream_t s);   // fills Buffers::row_span (height entries)
// counts, over every pixel classified "certain miss", the jitter positions (8x8 grid incl. both extremes) whose exact
// discriminant says HIT — must be 0; tested = positions evaluated (ptb_selftest_exact_math which = 3)
cudaError_t selftest_certain_miss(const FrameK &k, unsigned long long *mismatches, unsigned long long *tested, cudaStream_t s);
// out8 != nullptr: the launch also resolves every pixel it accumulates (accum / spp_total -> UNORM8, the arithmetic of
// resolve_kernel without the sRGB table) into out8; bgra selects the byte order
cudaError_t launch_csmain_accumulate(const FrameK &k, float4 *accum, uint32_t first_sample,
                                     uint32_t n_samples, uint32_t seed, bool jitter, cudaStream_t s,
                                     uint32_t *out8 = nullptr, float spp_total = 1.0f, bool bgra = false);
// the same for the row-span table of the 1 spp frame kernels: pixels of quads outside a row's span that HIT (must be 0)
cudaError_t selftest_row_span(const FrameK &k, const uint2 *row_span, unsigned long long *mismatches, unsigned long long *tested,
                              cudaStream_t s);
// selftest.cu — exhaustive proofs of the short exact sequences in ptb_math.cuh
//   which: 0 = sqrt_rn_fast over [2^-20,2^24), 1 = rcp_rn_fast over [2^-12,2^12),
//          2 = div_rn_uniform(v, D) for v in {0} U [2^-24, 2^17), 5 = the split-reciprocal form fma(v, dh, v*dl) likewise
cudaError_t selftest_exact(int which, float D, unsigned long long *mismatches, unsigned long long *tested,
                           cudaStream_t s);
// cached per (device, D); bit 0: the one-correction (3-instruction) division is exact for D, bit 1: so is the
// 2-instruction split-reciprocal form
int uniform_divisor_verified(float D, cudaStream_t s);
float reciprocal_low(float D);                               // RN(1/D - RN(1/D)), evaluated in binary64

// kernels_image.cu
cudaError_t image_tables_init();   // sRGB tables -> __constant__ (once per device)
cudaError_t launch_resolve(const float4 *accum, uint32_t *out8, uint32_t n_pixels, uint32_t spp_total,
                           uint32_t flags, cudaStream_t s);
cudaError_t launch_present(const uint32_t *src, uint32_t sw, uint32_t sh, uint32_t *dst, uint32_t dw,
                           uint32_t dh, uint32_t flags, cudaStream_t s);

// error: device word raised by a rendezvous that timed out (nullptr = none): the merge / clear then do nothing
cudaError_t launch_peer_merge(void *const *peer_accum, uint32_t n_peers, float4 *total, uint32_t *dst_img, uint32_t first,
                              uint32_t count, uint32_t spp_total, uint32_t flags, cudaStream_t s, const uint32_t *error = nullptr);
cudaError_t launch_accum_clear(float4 *accum, uint32_t n_pixels, const uint32_t *error, cudaStream_t s);

cudaError_t launch_peer_barrier(void *const *peer_flags, uint32_t n_peers, uint32_t self, uint32_t phase, uint32_t epoch,
                                uint32_t *error, cudaStream_t s);

// scene.cu — the scene path (n3-n7) behind ptb_compute_frame / ptb_render_samples / ptb_wait
void scene_destroy(ptb_context *ctx);
bool scene_resident(const ptb_context *ctx);
int scene_compute_frame(ptb_context *ctx);
int scene_render_samples(ptb_context *ctx, uint32_t first_sample, uint32_t n_samples, int jitter);
int scene_wait_check(ptb_context *ctx);
// api.cu — error plumbing shared by the translation units
int fail_msg(ptb_context *ctx, int code, const char *msg);
int fail_cuda(ptb_context *ctx, cudaError_t e, const char *what);

}  // namespace ptb

namespace ptb {
// Native multi-GPU frame loop (ptb_peer_group_* / ptb_peer_frame): sample-range sharding with the fused, pipelined
// peer-memory merge, driven entirely from this library (one C-ABI call per frame and rank).
struct PeerGroup {
    bool bound = false;
    cudaStream_t merge_stream = nullptr;          // highest priority: its small kernels must get SM slots under a render grid
    float4 *partial[2] = {nullptr, nullptr};      // frame k renders into partial[k & 1]; [0] is the context's own accumulation buffer
    uint32_t *image[2] = {nullptr, nullptr};      // frame k's merged image lands in image[k & 1] of the destination rank
    cudaEvent_t ev_rendered = nullptr;
    cudaEvent_t ev_released[2] = {nullptr, nullptr};   // partial[b] merged by every peer and cleared: free for frame k + 2
    cudaEvent_t ev_merged[2] = {nullptr, nullptr};     // image[b] complete (after the second rendezvous)
    bool released_pending[2] = {false, false};
    bool merged_valid[2] = {false, false};
    uint32_t n_peers = 0, self = 0, dst = 0;
    void *peer_partial[2][PTB_MAX_PEERS] = {};
    void *peer_flags[PTB_MAX_PEERS] = {};
    void *dst_image[2] = {nullptr, nullptr};
    uint32_t epoch[2] = {0, 0};
    uint64_t frame = 0;
};
}  // namespace ptb

struct ptb_context {
    ptb_config cfg{};
    ptb::FrameK k{};
    ptb::Buffers buf{};
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // asynchronous image read-back (ptb_read_rgba8_async): a copy stream ordered after the producer
    // by ev_img_ready; the next writer of the image waits on ev_copy_done.
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_img_ready = nullptr, ev_copy_done = nullptr;
    bool copy_pending = false;
    cudaStream_t copy_extra[3] = {nullptr, nullptr, nullptr};     // PTB200_COPY_SPLIT: further row bands of one read-back
    cudaEvent_t ev_copy_part[3] = {nullptr, nullptr, nullptr};
    // the same pair for buf.rgba8_alt (ptb_render_resolve swaps image, event and flag together)
    cudaEvent_t ev_copy_done_alt = nullptr;
    bool copy_pending_alt = false;
    bool timer_open = false;
    uint32_t accum_spp = 0;
    uint64_t launches = 0;
    std::string err;
    void *scene = nullptr;       // ptb::SceneState*, owned (scene.cu)
    // fused peer merge (ptb_peer_merge): this GPU's slab of the running total, and the IPC mappings it opened
    float4 *slab_total = nullptr;
    uint32_t slab_first = 0, slab_count = 0;
    std::vector<void *> ipc_opened;
    uint32_t *peer_flags = nullptr;   // 2 x PTB_MAX_PEERS epochs (arrive / done) + 1 error word, written by the peers
    ptb::PeerGroup *peer = nullptr;   // ptb_peer_group_create
};
