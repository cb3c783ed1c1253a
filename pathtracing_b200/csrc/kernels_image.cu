// kernels_image.cu — resolve ("tonemap") and present-equivalent kernels (HBM-bound byte work).
//
//   resolve : accum float4 / spp -> UNORM8 (optional sRGB OETF, optional BGRA order).
//             The UNORM8 clamp is what the reference's R8G8B8A8Unorm storage image implies
//             (RayTracing/Graphics/Renderer.cs:124); the OETF is what its preferred
//             B8G8R8A8Srgb swapchain format applies on write (Graphics/SwapChain.cs:154-165).
//   present : the arithmetic of the full-screen quad + PSMain
//             (Assets/Shaders/Source/Basic/PixelShader.hlsl:8-19; sampler nearest/repeat,
//             Renderer.cs:181-191; quad uv without flip, Renderer.cs:45-56) as a headless copy.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <mutex>

#include "ptb_internal.h"
#include "ptb_math.cuh"

namespace ptb {

// c_srgb_thresh[k] (k = 1..255) = smallest float whose sRGB-encoded UNORM8 value is >= k;
// encode(v) = #{k : v >= thresh[k]}.  c_srgb_lut8[i] = encode(i/255) for 8-bit sources.
__constant__ float c_srgb_thresh[256];
__constant__ uint8_t c_srgb_lut8[256];

static double srgb_eotf(double e) { return e <= 0.04045 ? e / 12.92 : std::pow((e + 0.055) / 1.055, 2.4); }

cudaError_t image_tables_init()
{
    float thresh[256];
    uint8_t lut[256];
    thresh[0] = 0.0f;
    for (int k = 1; k < 256; ++k) {
        const double b = srgb_eotf((k - 0.5) / 255.0);   // linear value where the code flips to k
        float f = (float)b;
        if ((double)f < b) f = std::nextafterf(f, 2.0f);
        thresh[k] = f;
    }
    for (int i = 0; i < 256; ++i) {
        const float v = (float)i / 255.0f;
        int code = 0;
        for (int k = 1; k < 256; ++k) code += (v >= thresh[k]);
        lut[i] = (uint8_t)code;
    }
    cudaError_t e = cudaMemcpyToSymbol(c_srgb_thresh, thresh, sizeof thresh);
    if (e != cudaSuccess) return e;
    return cudaMemcpyToSymbol(c_srgb_lut8, lut, sizeof lut);
}

// Branch-free 8-step binary search over the (shared-memory copy of the) threshold table.
__device__ __forceinline__ uint32_t srgb8_from_linear(const float *__restrict__ th, float v)
{
    if (!(v > 0.0f)) return 0u;   // NaN and negatives
    uint32_t lo = 0;              // invariant: v >= th[lo] (th[0] = 0)
#pragma unroll
    for (uint32_t step = 128; step > 0; step >>= 1) {
        const uint32_t mid = lo + step;   // <= 255
        lo = (v >= th[mid]) ? mid : lo;
    }
    return lo;
}

template <bool SRGB, bool BGRA>
__global__ void __launch_bounds__(256) resolve_kernel(const float4 *__restrict__ accum,
                                                      uint32_t *__restrict__ out8, uint32_t n_pixels,
                                                      float spp)
{
    __shared__ float th[256];
    if (SRGB) {
        th[threadIdx.x] = c_srgb_thresh[threadIdx.x];
        __syncthreads();
    }
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_pixels; i += stride) {
        const float4 a = __ldcs(accum + i);   // streamed once: evict-first
        const float r = fdiv(a.x, spp), g = fdiv(a.y, spp), b = fdiv(a.z, spp), al = fdiv(a.w, spp);
        uint32_t r8, g8, b8;
        if (SRGB) {
            r8 = srgb8_from_linear(th, r);
            g8 = srgb8_from_linear(th, g);
            b8 = srgb8_from_linear(th, b);
        } else {
            r8 = unorm8(r);
            g8 = unorm8(g);
            b8 = unorm8(b);
        }
        const uint32_t a8 = unorm8(al);
        out8[i] = BGRA ? (b8 | (g8 << 8) | (r8 << 16) | (a8 << 24)) : (r8 | (g8 << 8) | (b8 << 16) | (a8 << 24));
    }
}

cudaError_t launch_resolve(const float4 *accum, uint32_t *out8, uint32_t n_pixels, uint32_t spp_total,
                           uint32_t flags, cudaStream_t s)
{
    const int block = 256;
    int grid = (int)((n_pixels + block - 1) / block);
    const int cap = 148 * 8 * 4;   // a few waves of resident CTAs; grid-stride beyond that
    if (grid > cap) grid = cap;
    const float spp = (float)spp_total;
    const bool srgb = flags & PTB_RESOLVE_SRGB, bgra = flags & PTB_RESOLVE_BGRA;
    if (srgb && bgra) resolve_kernel<true, true><<<grid, block, 0, s>>>(accum, out8, n_pixels, spp);
    else if (srgb) resolve_kernel<true, false><<<grid, block, 0, s>>>(accum, out8, n_pixels, spp);
    else if (bgra) resolve_kernel<false, true><<<grid, block, 0, s>>>(accum, out8, n_pixels, spp);
    else resolve_kernel<false, false><<<grid, block, 0, s>>>(accum, out8, n_pixels, spp);
    return cudaGetLastError();
}

// ---- fused multi-GPU merge + resolve over NVLink peer memory (north_star stage n8) ----------------------------
// Sample-range sharding leaves G partial float4 accumulation buffers, one per GPU.  Instead of an NCCL reduce of the
// whole 133 MB frame to one GPU followed by a resolve there, every GPU runs THIS kernel over its own 1/G slab of the
// pixels: it loads the G partials of each pixel straight from the peers' memory (P2P loads through NVSwitch), adds them
// in rank order (a fixed order: the merged image does not depend on timing), folds the sum into its slab of the
// running total, resolves to UNORM8 and stores the pixel into the destination GPU's image (P2P store).  Per GPU
// (G-1)/G of one frame's partials cross NVLink inbound, spread over all G-1 peers at once through the switch, and 1/G of
// the RGBA8 image goes out; one launch, no staging buffer, no second pass for the resolve.  Each GPU clears its own
// partial buffer afterwards (local memset, ptb_accum_clear) instead of paying for remote zero stores.
struct PeerPtrs {
    float4 *p[PTB_MAX_PEERS];
};
struct PeerFlagPtrs {
    uint32_t *p[PTB_MAX_PEERS];
};

// NP = compile-time bound on the number of peers, UNROLL pixels per thread per iteration so that every thread keeps
// >= 8 independent 16 B loads in flight.  The grid is deliberately SMALL (one 256-thread block per SM): the kernel is
// bound by NVLink / HBM latency, not by threads, and a machine-filling grid on the high-priority merge stream would take
// every SM slot away from the next frame's render kernel for the whole duration of the merge — measured: no overlap at
// all with a 148x8-block grid (experiments/overlap_check.py, profiles/r01_notes.md).
template <int NP, int UNROLL>
__global__ void __launch_bounds__(256) peer_merge_kernel(const PeerPtrs peers, uint32_t n_peers, float4 *__restrict__ total,
                                                         uint32_t *__restrict__ dst_img, uint32_t first, uint32_t count, float spp,
                                                         uint32_t flags, const uint32_t *__restrict__ error)
{
    // a rendezvous that timed out (peer_barrier_kernel) means some peer's partials may be incomplete: merge nothing, the
    // error surfaces at the next ptb_wait and the partials stay where they are
    if (error && *reinterpret_cast<const volatile uint32_t *>(error)) return;
    __shared__ float th[256];
    const bool srgb = flags & PTB_RESOLVE_SRGB, bgra = flags & PTB_RESOLVE_BGRA;     // launch-uniform
    if (srgb) {
        th[threadIdx.x] = c_srgb_thresh[threadIdx.x];
        __syncthreads();
    }
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t base = blockIdx.x * blockDim.x + threadIdx.x; base < count; base += stride * UNROLL) {
        float4 parts[UNROLL][NP];
        float4 tot[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {                    // all loads in flight before the first add
            const uint32_t i = base + u * stride;
            if (i < count) {
#pragma unroll
                for (int q = 0; q < NP; ++q)
                    if ((uint32_t)q < n_peers) parts[u][q] = __ldcs(peers.p[q] + first + i);
                tot[u] = total[i];
            }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const uint32_t i = base + u * stride;
            if (i >= count) continue;
            float4 sum = parts[u][0];
#pragma unroll
            for (int q = 1; q < NP; ++q)
                if ((uint32_t)q < n_peers) {
                    sum.x = fadd(sum.x, parts[u][q].x); sum.y = fadd(sum.y, parts[u][q].y);
                    sum.z = fadd(sum.z, parts[u][q].z); sum.w = fadd(sum.w, parts[u][q].w);
                }
            float4 t = tot[u];
            t.x = fadd(t.x, sum.x); t.y = fadd(t.y, sum.y); t.z = fadd(t.z, sum.z); t.w = fadd(t.w, sum.w);
            total[i] = t;
            const float r = fdiv(t.x, spp), g = fdiv(t.y, spp), b = fdiv(t.z, spp), al = fdiv(t.w, spp);
            uint32_t r8, g8, b8;
            if (srgb) {
                r8 = srgb8_from_linear(th, r);
                g8 = srgb8_from_linear(th, g);
                b8 = srgb8_from_linear(th, b);
            } else {
                r8 = unorm8(r);
                g8 = unorm8(g);
                b8 = unorm8(b);
            }
            const uint32_t a8 = unorm8(al);
            dst_img[first + i] = bgra ? (b8 | (g8 << 8) | (r8 << 16) | (a8 << 24)) : (r8 | (g8 << 8) | (b8 << 16) | (a8 << 24));
        }
    }
}

cudaError_t launch_peer_merge(void *const *peer_accum, uint32_t n_peers, float4 *total, uint32_t *dst_img, uint32_t first,
                              uint32_t count, uint32_t spp_total, uint32_t flags, cudaStream_t s, const uint32_t *error)
{
    if (count == 0) return cudaSuccess;
    PeerPtrs pp{};
    for (uint32_t q = 0; q < n_peers; ++q) pp.p[q] = static_cast<float4 *>(peer_accum[q]);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // PTB_MERGE_BACKGROUND: one block per SM, so the merge shares the machine with a concurrent render kernel instead
    // of displacing it; otherwise a machine-filling grid (fastest when nothing else runs).  Measured on one B200, 4K frame,
    // 64-spp render 0.703 ms: merge behind it on the same stream +0.076 ms (8 blocks/SM) / +0.140 ms (1 block/SM); on a
    // second high-priority stream overlapping the NEXT render +0.061 ms (8/SM) / +0.041 ms (1/SM).
    static const int env_per_sm = [] {
        const char *e = std::getenv("PTB200_MERGE_BLOCKS_PER_SM");     // A/B experiments
        const int v = e ? std::atoi(e) : 0;
        return v >= 1 && v <= 8 ? v : 0;
    }();
    const int per_sm = env_per_sm ? env_per_sm : ((flags & PTB_MERGE_BACKGROUND) ? 1 : 8);
    const int block = 256;
    int grid = (int)((count + block - 1) / block);
    if (grid > sms * per_sm) grid = sms * per_sm;
    const float spp = (float)spp_total;
    if (n_peers <= 2) peer_merge_kernel<2, 4><<<grid, block, 0, s>>>(pp, n_peers, total, dst_img, first, count, spp, flags, error);
    else if (n_peers <= 4) peer_merge_kernel<4, 2><<<grid, block, 0, s>>>(pp, n_peers, total, dst_img, first, count, spp, flags, error);
    else if (n_peers <= 8) peer_merge_kernel<8, 1><<<grid, block, 0, s>>>(pp, n_peers, total, dst_img, first, count, spp, flags, error);
    else peer_merge_kernel<16, 1><<<grid, block, 0, s>>>(pp, n_peers, total, dst_img, first, count, spp, flags, error);
    return cudaGetLastError();
}

// Cross-GPU rendezvous through peer memory: one warp.  Lane q publishes `epoch` into slot [phase][self] of peer q's flag
// array with a system-scope RELEASE store (everything this GPU's stream did before — the render kernel's writes to its
// partial buffer — happens-before any peer's acquire that observes the flag), then spins with system-scope ACQUIRE loads
// until slot [phase][q] of its OWN array reaches `epoch`.  A 32-thread block fits into any SM slot that frees up, so the
// rendezvous is not starved by a large render grid running at lower priority — which is what happened to the
// (640-thread) NCCL kernels of a 1-element all-reduce in the pipelined merge (profiles/r01_notes.md).
// Bounded spin: after `timeout_ns` (default 10 s) the kernel gives up and raises *error (bit 1); the merge and clear
// kernels that follow on the stream read the word and do nothing, and ptb_wait reports PTB_ERR_STATE.
__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__global__ void peer_barrier_kernel(const PeerFlagPtrs peers, uint32_t n_peers, uint32_t self, uint32_t phase, uint32_t epoch,
                                    uint32_t *error, unsigned long long timeout_ns)
{
    const uint32_t q = threadIdx.x;
    if (q < n_peers) {
        st_release_sys(peers.p[q] + phase * PTB_MAX_PEERS + self, epoch);
        const uint32_t *mine = peers.p[self] + phase * PTB_MAX_PEERS + q;
        unsigned long long t0 = 0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > timeout_ns) {
                atomicOr(error, 2u);
                break;
            }
            __nanosleep(100);
        }
    }
    __syncwarp();
    __threadfence_system();      // the kernels that follow on this stream read peer memory: order them after every lane's acquire
}

// zero the partial accumulation buffer unless the rendezvous before it failed (see peer_merge_kernel)
__global__ void __launch_bounds__(256) accum_clear_kernel(float4 *__restrict__ accum, uint32_t n, const uint32_t *__restrict__ error)
{
    if (error && *reinterpret_cast<const volatile uint32_t *>(error)) return;
    const float4 z = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) accum[i] = z;
}

cudaError_t launch_accum_clear(float4 *accum, uint32_t n_pixels, const uint32_t *error, cudaStream_t s)
{
    if (n_pixels == 0) return cudaSuccess;
    const unsigned blocks = (unsigned)std::min<uint64_t>(((uint64_t)n_pixels + 255u) / 256u, 148u * 4u);
    accum_clear_kernel<<<blocks, 256, 0, s>>>(accum, n_pixels, error);
    return cudaGetLastError();
}

cudaError_t launch_peer_barrier(void *const *peer_flags, uint32_t n_peers, uint32_t self, uint32_t phase, uint32_t epoch,
                                uint32_t *error, cudaStream_t s)
{
    PeerFlagPtrs pf{};
    for (uint32_t q = 0; q < n_peers; ++q) pf.p[q] = static_cast<uint32_t *>(peer_flags[q]);
    static const unsigned long long timeout_ns = [] {
        const char *e = std::getenv("PTB200_BARRIER_TIMEOUT_MS");      // tests shorten the 10 s default
        const long long v = e ? std::atoll(e) : 0;
        return v > 0 ? (unsigned long long)v * 1000000ull : 10000000000ull;
    }();
    peer_barrier_kernel<<<1, 32, 0, s>>>(pf, n_peers, self, phase, epoch, error, timeout_ns);
    return cudaGetLastError();
}

template <bool SRGB>
__global__ void __launch_bounds__(256) present_kernel(const uint32_t *__restrict__ src, uint32_t sw, uint32_t sh,
                                                      uint32_t *__restrict__ dst, uint32_t dw, uint32_t dh)
{
    __shared__ uint8_t lut[256];
    if (SRGB) {
        lut[threadIdx.y * blockDim.x + threadIdx.x] = c_srgb_lut8[threadIdx.y * blockDim.x + threadIdx.x];
        __syncthreads();
    }
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t j = blockIdx.y * blockDim.y + threadIdx.y;
    if (i >= dw || j >= dh) return;
    // nearest texel of the pixel-centre uv: floor(((i+0.5)/dw)*sw) in exact integer form,
    // repeat addressing.
    const uint32_t sx = (uint32_t)(((2ull * i + 1ull) * sw) / (2ull * dw)) % sw;
    const uint32_t sy = (uint32_t)(((2ull * j + 1ull) * sh) / (2ull * dh)) % sh;
    const uint32_t p = __ldg(src + (size_t)sy * sw + sx);
    uint32_t r = p & 0xFFu, g = (p >> 8) & 0xFFu, b = (p >> 16) & 0xFFu;
    const uint32_t a = p >> 24;
    if (SRGB) {
        r = lut[r];
        g = lut[g];
        b = lut[b];
    }
    dst[(size_t)j * dw + i] = b | (g << 8) | (r << 16) | (a << 24);
}

cudaError_t launch_present(const uint32_t *src, uint32_t sw, uint32_t sh, uint32_t *dst, uint32_t dw,
                           uint32_t dh, uint32_t flags, cudaStream_t s)
{
    const dim3 block(32, 8);
    const dim3 grid((dw + 31) / 32, (dh + 7) / 8);
    if (flags & PTB_RESOLVE_SRGB) present_kernel<true><<<grid, block, 0, s>>>(src, sw, sh, dst, dw, dh);
    else present_kernel<false><<<grid, block, 0, s>>>(src, sw, sh, dst, dw, dh);
    return cudaGetLastError();
}

}  // namespace ptb
