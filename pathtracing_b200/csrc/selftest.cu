// selftest.cu — proofs by exhaustion of the short exactly-rounded sequences in ptb_math.cuh.
//
// The fast kernels may only replace an IEEE intrinsic by a shorter FMA sequence where the two
// agree on EVERY input the kernel can produce.  For sqrt and rcp the domain is a fixed float
// range, checked over all of its floats (ptb_selftest_exact_math, run by the -m gpu tests).
// For the launch-uniform screen divisor D (Test.hlsl:6 `resolution`) the one-correction
// division is checked against __fdiv_rn for every float v in {0} U [2^-24, 2^17) when a context
// with that D is first created (~0.3 ms on a B200, cached per device and divisor); if a single
// v disagrees the context silently uses the literal IEEE kernels instead.
#include <cstring>
#include <map>
#include <mutex>
#include <utility>

#include "ptb_internal.h"
#include "ptb_math.cuh"

namespace ptb {

struct SelfCounts {
    unsigned long long bad, n;
};

template <int WHICH>
__global__ void __launch_bounds__(256) selftest_kernel(uint32_t lo_bits, uint32_t hi_bits, float D, float Dh, float Dl,
                                                       SelfCounts *out)
{
    unsigned long long bad = 0, n = 0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t b = (uint64_t)lo_bits + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b < hi_bits; b += stride) {
        const float v = __uint_as_float((uint32_t)b);
        float got, ref;
        if (WHICH == 0) {
            got = sqrt_rn_fast(v);
            ref = __fsqrt_rn(v);
        } else if (WHICH == 1) {
            got = rcp_rn_fast(v);
            ref = __frcp_rn(v);
        } else if (WHICH == 2) {
            got = div_rn_uniform(v, D, Dh);
            ref = __fdiv_rn(v, D);
        } else {
            got = ffma(v, Dh, fmul(v, Dl));       // the split-reciprocal division of kernels_csmain.cu screen_uv_axes<true>
            ref = __fdiv_rn(v, D);
        }
        bad += (__float_as_uint(got) != __float_as_uint(ref));
        ++n;
    }
    // warp-reduce, one atomic per warp
    for (int o = 16; o > 0; o >>= 1) {
        bad += __shfl_down_sync(0xFFFFFFFFu, bad, o);
        n += __shfl_down_sync(0xFFFFFFFFu, n, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&out->bad, bad);
        atomicAdd(&out->n, n);
    }
}

static uint32_t f2u(float f)
{
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
}

float reciprocal_low(float D)
{
    const float dh = 1.0f / D;
    return (float)(1.0 / (double)D - (double)dh);
}

cudaError_t selftest_exact(int which, float D, unsigned long long *mismatches, unsigned long long *tested,
                           cudaStream_t s)
{
    SelfCounts *d = nullptr, h{0, 0};
    cudaError_t e = cudaMalloc(&d, sizeof *d);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(d, 0, sizeof *d, s);
    const int grid = 148 * 8, block = 256;
    if (e == cudaSuccess) {
        if (which == 0) {
            selftest_kernel<0><<<grid, block, 0, s>>>(f2u(9.5367431640625e-7f) /*2^-20*/, f2u(16777216.0f), 0.f, 0.f, 0.f, d);
        } else if (which == 1) {
            selftest_kernel<1><<<grid, block, 0, s>>>(f2u(2.44140625e-4f) /*2^-12*/, f2u(4096.0f), 0.f, 0.f, 0.f, d);
        } else {
            const float Dh = 1.0f / D;   // RN(1/D): IEEE binary32 division on the host == __frcp_rn(D) on the device
            // v = 0 is covered analytically (0*Dh = 0, fma(-0,D,0) = 0, fma(0,Dh,0) = 0 == 0/D for finite D != 0)
            const float Dl = reciprocal_low(D);
            if (which == 2) selftest_kernel<2><<<grid, block, 0, s>>>(f2u(5.9604644775390625e-8f) /*2^-24*/, f2u(131072.0f), D, Dh, Dl, d);
            else selftest_kernel<5><<<grid, block, 0, s>>>(f2u(5.9604644775390625e-8f) /*2^-24*/, f2u(131072.0f), D, Dh, Dl, d);
        }
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(&h, d, sizeof h, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(d);
    if (mismatches) *mismatches = h.bad;
    if (tested) *tested = h.n;
    return e;
}

int uniform_divisor_verified(float D, cudaStream_t s)
{
    static std::mutex mu;
    static std::map<std::pair<int, uint32_t>, int> cache;
    // the sequences are sign-symmetric and exponent-shift invariant only away from under/overflow,
    // so each distinct D in the supported magnitude window is verified on its own.
    const float a = D < 0 ? -D : D;
    if (!(a >= 1.0f / 1048576.0f && a <= 1048576.0f)) return 0;   // also rejects NaN/inf/0
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    std::lock_guard<std::mutex> lock(mu);
    const auto key = std::make_pair(dev, f2u(D));
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
    unsigned long long bad = 1, n = 0;
    int ok = (selftest_exact(2, D, &bad, &n, s) == cudaSuccess && bad == 0 && n > 0) ? 1 : 0;
    bad = 1;
    n = 0;
    if (ok && selftest_exact(5, D, &bad, &n, s) == cudaSuccess && bad == 0 && n > 0) ok |= 2;
    cache[key] = ok;
    return ok;
}

}  // namespace ptb
