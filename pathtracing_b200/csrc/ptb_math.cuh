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

// ---- short exactly-rounded sequences ---------------------------------------------------------
// Each is bit-identical to the IEEE intrinsic it replaces on the stated domain; the domains are
// proven by exhaustion on the GPU (selftest.cu / experiments/exactmath.cu) and the callers
// guarantee them (launch-uniform checks in api.cu, per-sample exponent guards in the kernels).

__device__ __forceinline__ float mufu_rsq(float x)
{
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float mufu_rcp(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// == __fsqrt_rn(x) for every float x in [2^-20, 2^24) (exhaustive).
__device__ __forceinline__ float sqrt_rn_fast(float x)
{
    const float r = mufu_rsq(x);
    const float s = fmul(x, r), h = fmul(r, 0.5f);
    return ffma(ffma(-s, s, x), h, s);
}

// == __frcp_rn(y) for every float y in [2^-12, 2^12) (exhaustive).
__device__ __forceinline__ float rcp_rn_fast(float y)
{
    const float q = mufu_rcp(y);
    return ffma(ffma(-y, q, 1.0f), q, q);
}

// v / y given yh = RN(1/y): q0 = RN(v*yh) (always within 1 ulp), then ONE Markstein correction.
// Correctly rounded for EVERY pair of binary32 significands — proven by exhaustion over all 2^46
// pairs on a B200 (experiments/div3_exhaustive.cu, 57.5 s; profiles/r01_div3_exhaustive.txt), and
// exact scaling by powers of two extends that to all normal operands whose residual v - q0*y does
// not underflow (|v| >= 2^-100 or v == 0; the callers' operands are screen/unit-sphere sized).
__device__ __forceinline__ float div_rn_by_rcp(float v, float y, float yh)
{
    const float q0 = fmul(v, yh);
    return ffma(ffma(-q0, y, v), yh, q0);
}

// v / D for a launch-uniform D with Dh = RN(1/D): the same one-correction division; additionally
// re-verified for this D over every float v in [2^-24, 2^17) when the context is created (selftest.cu).
__device__ __forceinline__ float div_rn_uniform(float v, float D, float Dh)
{
    const float q0 = fmul(v, Dh);
    return ffma(ffma(-q0, D, v), Dh, q0);
}

// true iff 2^-20 <= x < 2^24 (x a non-negative finite float): one IADD + one ISETP.
__device__ __forceinline__ bool in_fast_range(float x)
{
    return (__float_as_uint(x) - 0x35800000u) < (0x4B800000u - 0x35800000u);
}

// ---- RNG: counter-based pcg3d (Jarzynski & Olano, "Hash Functions for GPU Rendering", JCGT 2020)
// keyed (pixel index, sample index, seed).  No reference counterpart (Test.hlsl is deterministic);
// this is the frozen spec shared with the CPU oracle.
struct U3 {
    uint32_t x, y, z;
};

__host__ __device__ __forceinline__ U3 pcg3d(uint32_t x, uint32_t y, uint32_t z)
{
    x = x * 1664525u + 1013904223u;
    y = y * 1664525u + 1013904223u;
    z = z * 1664525u + 1013904223u;
    x += y * z;
    y += z * x;
    z += x * y;
    x ^= x >> 16;
    y ^= y >> 16;
    z ^= z >> 16;
    x += y * z;
    y += z * x;
    z += x * y;
    return U3{x, y, z};
}

// 23 random bits -> float in [1,2) (exact); the callers subtract 1 or fold it into an add.
__device__ __forceinline__ float bits_to_unit_plus_one(uint32_t v) { return __uint_as_float((v >> 9) | 0x3F800000u); }
// uniform in [0,1): (v >> 9) * 2^-23 (exact).
__device__ __forceinline__ float bits_to_unit(uint32_t v) { return fsub(bits_to_unit_plus_one(v), 1.0f); }

}  // namespace ptb

// ---- packed FP32x2 forms (Blackwell FFMA2 / FMUL2 / FADD2: two IEEE-rounded lanes per issue slot)
// The CSMain kernels are issue-bound, so evaluating two samples per instruction halves the FP
// issue pressure without changing any lane's rounding (each lane is an independent *_rn op).
namespace ptb {

typedef float2 F2;

__device__ __forceinline__ F2 splat(float s) { return make_float2(s, s); }
__device__ __forceinline__ F2 fmul2(F2 a, F2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ F2 fadd2(F2 a, F2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ F2 ffma2(F2 a, F2 b, F2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ F2 neg2(F2 a) { return make_float2(-a.x, -a.y); }
// A multiply whose result feeds an ADD.  ptxas (12.9) contracts mul.rn.f32x2 + add.rn.f32x2 into one
// FFMA2 — unlike the scalar *_rn forms it does not honour .rn as "do not fuse", even with -fmad=false,
// and it also folds fma(a,b,-0.0) back to a multiply first.  An FMA whose addend is a RUN-TIME -0.0
// (FrameK::nzero) is opaque to it and rounds exactly like the multiply: RN(a*b + (-0)) == RN(a*b),
// signed zeros included.  Same issue cost as FMUL2.
__device__ __forceinline__ F2 fmul2a(F2 a, F2 b, F2 nzero) { return __ffma2_rn(a, b, nzero); }

// Returns -sqrt(x) given xn = -x (x in [2^-20, 2^24)): the negated length is what every later
// correction step consumes (fma(q, -len, v)), so no separate negations are issued.
//   s = x*r, e = x - s*s, sqrt = s + e*(r/2)   ==>   -sqrt = fma(sn*sn - x, r/2, sn), sn = -s
__device__ __forceinline__ F2 neg_sqrt_rn_fast2(F2 xn)
{
    const F2 r = make_float2(mufu_rsq(-xn.x), mufu_rsq(-xn.y));
    const F2 sn = fmul2(xn, r), h = fmul2(r, splat(0.5f));
    return ffma2(ffma2(sn, sn, xn), h, sn);
}

// RN(1/y) given yn = -y (y in [2^-12, 2^12)).
__device__ __forceinline__ F2 rcp_rn_fast2_neg(F2 yn)
{
    const F2 q = make_float2(mufu_rcp(-yn.x), mufu_rcp(-yn.y));
    return ffma2(ffma2(yn, q, splat(1.0f)), q, q);
}

// v / y given yn = -y and yh = RN(1/y): one Markstein correction (see div_rn_by_rcp).
__device__ __forceinline__ F2 div_rn_by_rcp2(F2 v, F2 yn, F2 yh)
{
    const F2 q0 = fmul2(v, yh);
    return ffma2(ffma2(q0, yn, v), yh, q0);
}

// v / D for the launch-uniform D given Dn = -D and Dh = RN(1/D): one correction (see div_rn_uniform).
__device__ __forceinline__ F2 div_rn_uniform2(F2 v, F2 Dn, F2 Dh)
{
    const F2 q0 = fmul2(v, Dh);
    return ffma2(ffma2(q0, Dn, v), Dh, q0);
}

}  // namespace ptb
