// kernels_csmain.cu — sm_100a kernels for the reference's per-pixel tracing path.
//
//   csmain_frame_*      : one launch == one Renderer.ComputeFrame dispatch
//                         (RayTracing/Graphics/Renderer.cs:1006-1040) of CSMain
//                         (RayTracing/Assets/Shaders/Source/Ray/Test.hlsl:3-40), 1 primary ray per
//                         pixel, stored as packed R8G8B8A8 UNORM (Renderer.cs:124) and optionally as
//                         the shader-side float4 and a hit-id plane.
//   csmain_accumulate   : north_star stage n1/n2 — spp jittered samples per pixel summed
//                         in-kernel into a float4 accumulation buffer.
//
// Arithmetic is the frozen "one rounding per op" evaluation of Test.hlsl (ptb_math.cuh), so
// results are bit-identical to the CPU oracle's.  These kernels are FP32-issue bound (multi-spp)
// or HBM-write / launch bound (1 spp); there is no dense contraction, so no tensor cores.
#include "ptb_internal.h"
#include "ptb_math.cuh"

namespace ptb {

// Launch-uniform terms of Test.hlsl:16,19 (oc, c) evaluated once per thread.
struct SphereTerms {
    V3 cam, centre, oc;
    float c;
};

__device__ __forceinline__ SphereTerms sphere_terms(const FrameK &k)
{
    SphereTerms s;
    s.cam = V3{k.cam[0], k.cam[1], k.cam[2]};
    s.centre = V3{k.centre[0], k.centre[1], k.centre[2]};
    s.oc = V3{fsub(s.cam.x, s.centre.x), fsub(s.cam.y, s.centre.y), fsub(s.cam.z, s.centre.z)};  // :16
    s.c = fsub(dot3(s.oc, s.oc), fmul(k.radius, k.radius));                                       // :19
    return s;
}

// Test.hlsl:7 for one axis: uv = (id / resolution) * 2 - 1.
__device__ __forceinline__ float screen_uv(float p, float div)
{
    return fsub(fmul(fdiv(p, div), 2.0f), 1.0f);
}

// Test.hlsl:10-37 given uv.  Returns true on hit (disc > 0, strict).
__device__ __forceinline__ bool csmain_shade(const SphereTerms &s, float uvx, float uvy, float4 &col)
{
    const V3 d = normalize3(V3{uvx, uvy, -1.0f});                 // :10
    const float a = dot3(d, d);                                   // :17
    const float b = fmul(2.0f, dot3(s.oc, d));                    // :18
    const float disc = fsub(fmul(b, b), fmul(fmul(4.0f, a), s.c)); // :21
    if (disc > 0.0f) {                                            // :24
        const float t = fdiv(fsub(-b, fsqrt(disc)), fmul(2.0f, a));  // :27
        const V3 P{fadd(s.cam.x, fmul(t, d.x)), fadd(s.cam.y, fmul(t, d.y)), fadd(s.cam.z, fmul(t, d.z))};  // :28
        const V3 n = normalize3(V3{fsub(P.x, s.centre.x), fsub(P.y, s.centre.y), fsub(P.z, s.centre.z)});   // :29
        col = make_float4(fadd(fmul(n.x, 0.5f), 0.5f), fadd(fmul(n.y, 0.5f), 0.5f),
                          fadd(fmul(n.z, 0.5f), 0.5f), 1.0f);      // :31
        return true;
    }
    col = make_float4(uvx, uvy, 0.0f, 1.0f);                      // :36
    return false;
}

// ---- 1 spp frame, 4 pixels per thread, one STG.128 of packed RGBA8 per thread -------------
// Requires width % 4 == 0 (16 B aligned rows).  Block (32,8): a warp covers 128 consecutive
// pixels of one row = 512 contiguous bytes of RGBA8.
template <bool KEEP_F32, bool KEEP_HIT>
__global__ void __launch_bounds__(256) csmain_frame_vec4(const FrameK k, uint4 *__restrict__ rgba8,
                                                         float4 *__restrict__ f32, int4 *__restrict__ hit)
{
    const uint32_t xq = blockIdx.x * blockDim.x + threadIdx.x;   // quad index along x
    const uint32_t y = blockIdx.y * blockDim.y + threadIdx.y;
    const uint32_t wq = k.width >> 2;
    if (xq >= wq || y >= k.height) return;                       // the guard Test.hlsl:39 lacks

    const SphereTerms s = sphere_terms(k);
    const float uvy = screen_uv(__uint2float_rn(y), k.div_y);
    uint32_t px8[4];
    int32_t hid[4];
    const size_t q = (size_t)y * wq + xq;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float uvx = screen_uv(__uint2float_rn(xq * 4u + i), k.div_x);
        float4 col;
        const bool h = csmain_shade(s, uvx, uvy, col);
        px8[i] = pack_rgba8(col.x, col.y, col.z, col.w);
        hid[i] = h ? 0 : -1;
        if (KEEP_F32) f32[q * 4 + i] = col;
    }
    rgba8[q] = make_uint4(px8[0], px8[1], px8[2], px8[3]);
    if (KEEP_HIT) hit[q] = make_int4(hid[0], hid[1], hid[2], hid[3]);
}

// ---- 1 spp frame, 1 pixel per thread (any width) ------------------------------------------
template <bool KEEP_F32, bool KEEP_HIT>
__global__ void __launch_bounds__(256) csmain_frame_scalar(const FrameK k, uint32_t *__restrict__ rgba8,
                                                           float4 *__restrict__ f32, int32_t *__restrict__ hit)
{
    const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= k.width || y >= k.height) return;
    const SphereTerms s = sphere_terms(k);
    float4 col;
    const bool h = csmain_shade(s, screen_uv(__uint2float_rn(x), k.div_x),
                                screen_uv(__uint2float_rn(y), k.div_y), col);
    const size_t i = (size_t)y * k.width + x;
    rgba8[i] = pack_rgba8(col.x, col.y, col.z, col.w);
    if (KEEP_F32) f32[i] = col;
    if (KEEP_HIT) hit[i] = h ? 0 : -1;
}

template <bool F, bool H>
static cudaError_t launch_frame_t(const FrameK &k, const Buffers &b, cudaStream_t s)
{
    const dim3 block(32, 8);
    if ((k.width & 3u) == 0) {
        const dim3 grid(((k.width >> 2) + 31) / 32, (k.height + 7) / 8);
        csmain_frame_vec4<F, H><<<grid, block, 0, s>>>(k, reinterpret_cast<uint4 *>(b.rgba8), b.f32,
                                                       reinterpret_cast<int4 *>(b.hit));
    } else {
        const dim3 grid((k.width + 31) / 32, (k.height + 7) / 8);
        csmain_frame_scalar<F, H><<<grid, block, 0, s>>>(k, b.rgba8, b.f32, b.hit);
    }
    return cudaGetLastError();
}

cudaError_t launch_csmain_frame(const FrameK &k, const Buffers &b, cudaStream_t s)
{
    const bool f = b.f32 != nullptr, h = b.hit != nullptr;
    if (f && h) return launch_frame_t<true, true>(k, b, s);
    if (f) return launch_frame_t<true, false>(k, b, s);
    if (h) return launch_frame_t<false, true>(k, b, s);
    return launch_frame_t<false, false>(k, b, s);
}

// ---- multi-sample accumulation -------------------------------------------------------------
// One thread per pixel, samples looped in-kernel: the float4 read-modify-write of the
// accumulation buffer (32 B/pixel/launch) is amortised over n_samples, leaving the kernel
// FP32-issue bound.  Block (32,8) keeps a warp on one row (coalesced 512 B accum access) and a
// block on a 32x8 tile (hit/miss is spatially coherent, so few warps diverge).
template <bool JITTER>
__global__ void __launch_bounds__(256) csmain_accumulate(const FrameK k, float4 *__restrict__ accum,
                                                         uint32_t first_sample, uint32_t n_samples,
                                                         uint32_t seed_hash)
{
    const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= k.width || y >= k.height) return;
    const SphereTerms s = sphere_terms(k);
    const uint32_t pix = y * k.width + x;
    const float fx = __uint2float_rn(x), fy = __uint2float_rn(y);

    float4 sum = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    for (uint32_t i = 0; i < n_samples; ++i) {
        float px = fx, py = fy;
        if (JITTER) {
            uint32_t st = rng_init(pix, first_sample + i, seed_hash);
            px = fadd(fx, rng_float(st));
            py = fadd(fy, rng_float(st));
        }
        float4 col;
        csmain_shade(s, screen_uv(px, k.div_x), screen_uv(py, k.div_y), col);
        sum.x = fadd(sum.x, col.x);
        sum.y = fadd(sum.y, col.y);
        sum.z = fadd(sum.z, col.z);
        sum.w = fadd(sum.w, col.w);
    }
    float4 acc = accum[pix];
    acc.x = fadd(acc.x, sum.x);
    acc.y = fadd(acc.y, sum.y);
    acc.z = fadd(acc.z, sum.z);
    acc.w = fadd(acc.w, sum.w);
    accum[pix] = acc;
}

cudaError_t launch_csmain_accumulate(const FrameK &k, float4 *accum, uint32_t first_sample,
                                     uint32_t n_samples, uint32_t seed, bool jitter, cudaStream_t s)
{
    const dim3 block(32, 8);
    const dim3 grid((k.width + 31) / 32, (k.height + 7) / 8);
    const uint32_t seed_hash = pcg_hash(seed);
    if (jitter)
        csmain_accumulate<true><<<grid, block, 0, s>>>(k, accum, first_sample, n_samples, seed_hash);
    else
        csmain_accumulate<false><<<grid, block, 0, s>>>(k, accum, first_sample, n_samples, seed_hash);
    return cudaGetLastError();
}

}  // namespace ptb
