// ptb_math.cuh — exactly-rounded float32 helpers and the counter-based RNG for the sm_100a
// kernels.  Every arithmetic helper maps to ONE IEEE-754 binary32 round-to-nearest-even
// operation (the *_rn intrinsics are never contracted into FMAs by nvcc), which is what makes
// the kernels bit-identical to a CPU evaluation of the same expression tree.
//
// Arithmetic spec of the reference path: RayTracing/Assets/Shaders/Source/Ray/Test.hlsl:6-37,
// frozen as "every op rounded once, no contraction, left-to-right" (DESIGN.md §3).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace ptb {

__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float fsqrt(float a) { return __fsqrt_rn(a); }
__device__ __forceinline__ float ffma(float a, float b, float c) { return __fmaf_rn(a, b, c); }

struct V3 {
    float x, y, z;
};

// dot(a,b) = ((a0*b0 + a1*b1) + a2*b2), one rounding per op.
__device__ __forceinline__ float dot3(V3 a, V3 b)
{
    return fadd(fadd(fmul(a.x, b.x), fmul(a.y, b.y)), fmul(a.z, b.z));
}

// normalize(v) = v / sqrt(dot(v,v)), three correctly-rounded divisions.
__device__ __forceinline__ V3 normalize3(V3 v)
{
    const float len = fsqrt(dot3(v, v));
    return V3{fdiv(v.x, len), fdiv(v.y, len), fdiv(v.z, len)};
}

// f32 -> UNORM8 store conversion (the effective store of Test.hlsl:39 into the
// R8G8B8A8Unorm image of Renderer.cs:124): NaN -> 0, clamp, *255, round-nearest-even.
__device__ __forceinline__ uint32_t unorm8(float v)
{
    return __float2uint_rn(fmul(__saturatef(v), 255.0f));
}

__device__ __forceinline__ uint32_t pack_rgba8(float r, float g, float b, float a)
{
    return unorm8(r) | (unorm8(g) << 8) | (unorm8(b) << 16) | (unorm8(a) << 24);
}

// ---- RNG: PCG hash keyed on (pixel, sample, seed), PCG-RXS-M-XS-32 stream ----------------
__host__ __device__ __forceinline__ uint32_t pcg_hash(uint32_t v)
{
    const uint32_t state = v * 747796405u + 2891336453u;
    const uint32_t word = ((state >> ((state >> 28) + 4u)) ^ state) * 277803737u;
    return (word >> 22) ^ word;
}

__host__ __device__ __forceinline__ uint32_t rng_init(uint32_t pixel, uint32_t sample, uint32_t seed_hash)
{
    // seed_hash = pcg_hash(seed) is launch-uniform and hoisted by the callers.
    return pcg_hash(pixel + pcg_hash(sample + seed_hash));
}

__device__ __forceinline__ uint32_t rng_next(uint32_t &state)
{
    state = state * 747796405u + 2891336453u;
    const uint32_t word = ((state >> ((state >> 28) + 4u)) ^ state) * 277803737u;
    return (word >> 22) ^ word;
}

// uniform in [0,1): 24 random bits scaled by 2^-24 (exact).
__device__ __forceinline__ float rng_float(uint32_t &state)
{
    return fmul(__uint2float_rn(rng_next(state) >> 8), 5.9604644775390625e-8f);
}

}  // namespace ptb
